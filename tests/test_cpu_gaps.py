"""Round-2 gap closers on the host side: the non-IID listing on real PNG folders
(/root/reference/fed_model.py:161-164), the on-disk CIFAR-10 reader behind the dense script
(/root/reference/dist_model_tf_dense.py:120-121), early-stopped prefetch pipelines, unambiguous Keras
weight-name resolution and the extra command-line overrides of SURVEY.md 5.6."""
import pickle
import threading
import time

import numpy as np
import pytest
import torch

from idc_models_b200 import workloads as W
from idc_models_b200.data import cifar
from idc_models_b200.data.dataset import Dataset


def _write_pngs(root, per_class=6, size=12):
    from PIL import Image
    rng = np.random.default_rng(0)
    for cls in ("0", "1"):
        d = root / "data" / "balanced_IDC_30k" / cls
        d.mkdir(parents=True)
        for i in range(per_class):
            img = rng.integers(0, 255, (size, size, 3), dtype=np.uint8)
            if cls == "1":
                img[..., 0] = 255               # recognisable class marker
            Image.fromarray(img).save(d / f"p{i}_class{cls}.png")


def test_class_sorted_listing_on_real_folders_puts_class_1_before_class_0(tmp_path):
    _write_pngs(tmp_path)
    ds, n = W._idc_dataset(str(tmp_path), "data/balanced_IDC_30k/*/*", 0, (10, 10, 3), False, seed=3,
                           label_as_int=True, order="class_sorted")
    labels = [int(y) for _, y in ds]
    assert n == 12 and labels == [1] * 6 + [0] * 6
    # two iterations give the same frozen order (client shards must not move between rounds)
    assert [int(y) for _, y in ds] == labels
    # and the IID listing mixes the classes
    ds2, _ = W._idc_dataset(str(tmp_path), "data/balanced_IDC_30k/*/*", 0, (10, 10, 3), False, seed=3,
                            label_as_int=True, order="shuffled")
    mixed = [int(y) for _, y in ds2]
    assert sorted(mixed) == [0] * 6 + [1] * 6 and mixed != labels
    x, _ = next(iter(ds))
    assert np.asarray(x).shape == (10, 10, 3) and np.asarray(x)[..., 0].min() > 0.99   # a class-1 patch, resized


def _fake_cifar_planes(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)
    y = rng.integers(0, 10, n).astype(np.int64)
    planes = x.transpose(0, 3, 1, 2).reshape(n, 3072)
    return x, y, planes


@pytest.mark.parametrize("layout", ["bin", "py"])
def test_cifar10_reader_reads_both_distribution_layouts(tmp_path, layout):
    d = tmp_path / "data" / ("cifar-10-batches-" + layout)
    d.mkdir(parents=True)
    want_x, want_y = [], []
    for i, name in enumerate(["data_batch_1", "data_batch_2", "test_batch"]):
        x, y, planes = _fake_cifar_planes(5, i)
        if layout == "bin":
            rec = np.concatenate([y.astype(np.uint8)[:, None], planes], axis=1)
            rec.tofile(d / (name + ".bin"))
        else:
            with open(d / name, "wb") as fh:
                pickle.dump({b"data": planes, b"labels": y.tolist()}, fh)
        want_x.append(x)
        want_y.append(y)
    found = cifar.find_cifar10(str(tmp_path))
    assert found == str(d)
    (xtr, ytr), (xte, yte) = cifar.load_cifar10(found)
    assert xtr.shape == (10, 32, 32, 3) and xtr.dtype == np.uint8
    assert np.array_equal(xtr, np.concatenate(want_x[:2])) and np.array_equal(ytr, np.concatenate(want_y[:2]))
    assert np.array_equal(xte, want_x[2]) and np.array_equal(yte, want_y[2])
    # the dense workload picks it up ahead of the synthetic stand-in
    train, test, n = W._cifar10_datasets(str(tmp_path), 123, False, 0)
    assert n == 10 and len(train) == 10 and len(test) == 5
    img, lab = next(iter(train))
    assert np.array_equal(np.asarray(img), want_x[0][0]) and int(lab) == int(want_y[0][0])
    # --synthetic still forces the stand-in
    _, _, n2 = W._cifar10_datasets(str(tmp_path), 123, True, 0)
    assert n2 == 123


def test_prefetch_producer_stops_when_the_consumer_leaves_early():
    pulled = []

    def gen():
        for i in range(10_000):
            pulled.append(i)
            yield i

    ds = Dataset.from_generator(gen, 10_000).prefetch(2)
    before = threading.active_count()
    it = iter(ds)
    assert [next(it) for _ in range(3)] == [0, 1, 2]
    it.close()                                   # what a `break` out of a for loop does
    deadline = time.time() + 5
    while threading.active_count() > before and time.time() < deadline:
        time.sleep(0.02)
    assert threading.active_count() <= before, "the prefetch thread is still alive"
    assert len(pulled) < 20                      # it did not run ahead through the whole source either
    # a fresh iteration still works from the start
    assert next(iter(ds)) == 0


def test_keras_h5_short_weight_names_resolve_to_the_shallowest_match(tmp_path):
    """DenseNet's final BatchNorm is called 'bn': 'bn/gamma' also ends 'conv1/bn/gamma'. The loader must not
    depend on the order in which the file lists its groups."""
    from idc_models_b200.io import keras_h5

    class Net:
        def __init__(self):
            self.got = None

        def weight_names(self):
            return ["bn/gamma", "conv1/bn/gamma"]

        def set_named_weights(self, found, strict=False):
            self.got = found
            return [n for n in self.weight_names() if n not in found]

    deep = np.full(64, 2.0, np.float32)
    shallow = np.full(1024, 3.0, np.float32)
    for order in (("densenet121/conv1/bn/gamma:0", "densenet121/bn/gamma:0"),
                  ("densenet121/bn/gamma:0", "densenet121/conv1/bn/gamma:0")):
        arrays = {"densenet121/conv1/bn/gamma:0": deep, "densenet121/bn/gamma:0": shallow}
        path = str(tmp_path / "w.h5")
        keras_h5.write_h5(path, {k: arrays[k] for k in order}, [])
        net = Net()
        assert keras_h5.load_into(net, path) == []
        assert net.got["bn/gamma"].shape == (1024,) and net.got["conv1/bn/gamma"].shape == (64,)


def test_every_appendix_b_constant_has_a_flag():
    """dtype, GPU count, shuffle buffer, split fractions and the federated validation_steps."""
    import argparse
    ap = argparse.ArgumentParser()
    W._common_flags(ap)
    a = ap.parse_args(["--dtype", "fp32", "--gpus", "1", "--shuffle-buffer", "64", "--train-split", "0.7",
                       "--val-split", "0.15"])
    assert (a.dtype, a.gpus, a.shuffle_buffer, a.train_split, a.val_split) == ("fp32", 1, 64, 0.7, 0.15)
    assert W._executor_kind(a, "cuda") == "reference"
    with pytest.raises(SystemExit):
        a.gpus = 4
        W._check_gpus(a, type("Wd", (), {"size": 1})())


def test_mobile_script_honours_dtype_and_split_flags(tmp_path):
    out = W.main_transfer(W.MOBILE, [str(tmp_path), "--synthetic", "--dataset-size", "96", "--batch-size", "8",
                                     "--initial-epochs", "1", "--fine-tune-epochs", "1", "--steps-per-epoch", "2",
                                     "--validation-steps", "1", "--device", "cpu", "--dtype", "fp32",
                                     "--train-split", "0.5", "--val-split", "0.25", "--shuffle-buffer", "16"])
    assert type(out["model"].executor).__name__ == "RefExecutor"
    assert np.isfinite(out["history_fine"].history["loss"][-1])


def test_cifar_python_batches_cannot_name_arbitrary_callables(tmp_path):
    """a 'dataset file' whose pickle refers to os.system (or anything else a batch does not need) is refused"""
    import pickle
    d = tmp_path / "data" / "cifar-10-batches-py"
    d.mkdir(parents=True)

    class Evil:
        def __reduce__(self):
            import os
            return (os.system, ("echo pwned > /dev/null",))
    for name in ["data_batch_%d" % i for i in range(1, 6)] + ["test_batch"]:
        with open(d / name, "wb") as f:
            pickle.dump({b"data": Evil(), b"labels": [0]}, f, protocol=2)
    with pytest.raises(pickle.UnpicklingError, match="not something a dataset file needs"):
        cifar.load_cifar10(str(d))
