"""CPU unit tests: optimizers, losses, metrics, History bookkeeping, dataset ops."""
import math

import numpy as np
import pytest
import torch

from idc_models_b200.data.dataset import Dataset, prepare_for_training
from idc_models_b200.data import synthetic
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim import optimizers as O
from idc_models_b200.train import losses as L
from idc_models_b200.train import metrics as M


class _Store:
    def __init__(self, w):
        self.master = w.clone()
        self.grad = torch.zeros_like(w)
        self.shadow = None


def test_rmsprop_matches_keras_closed_form():
    w0 = torch.tensor([1.0, -2.0, 0.5, 3.0])
    st = _Store(w0)
    opt = O.RMSprop(learning_rate=0.1).bind(st, [(0, 4)])
    g = torch.tensor([0.5, -1.0, 2.0, 0.0])
    ms = torch.zeros(4)
    w = w0.clone()
    for _ in range(3):
        st.grad.copy_(g)
        opt.apply()
        ms = 0.9 * ms + 0.1 * g * g
        w = w - 0.1 * g / (ms.sqrt() + 1e-7)
    assert torch.allclose(st.master, w, atol=1e-6)
    assert st.grad.abs().max() == 0           # apply() re-zeroes the bucket


def test_sgd_and_adam_closed_form():
    w0 = torch.tensor([1.0, -2.0, 0.5, 3.0])
    g = torch.tensor([0.5, -1.0, 2.0, 0.25])
    st = _Store(w0)
    opt = O.SGD(learning_rate=1.0).bind(st, [(0, 4)])
    st.grad.copy_(g)
    opt.apply(grad_scale=0.5)
    assert torch.allclose(st.master, w0 - 0.5 * g)
    st = _Store(w0)
    opt = O.Adam(learning_rate=0.01).bind(st, [(0, 4)])
    m = v = torch.zeros(4)
    w = w0.clone()
    for t in range(1, 4):
        st.grad.copy_(g)
        opt.apply()
        m = 0.9 * m + 0.1 * g
        v = 0.999 * v + 0.001 * g * g
        lr_t = 0.01 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        w = w - lr_t * m / (v.sqrt() + 1e-7)
    assert torch.allclose(st.master, w, atol=1e-6)


def test_frozen_ranges_are_untouched():
    w0 = torch.arange(256, dtype=torch.float32)
    st = _Store(w0)
    opt = O.RMSprop(0.1).bind(st, [(64, 128)])
    st.grad.fill_(1.0)
    opt.apply()
    assert torch.equal(st.master[:64], w0[:64]) and torch.equal(st.master[128:], w0[128:])
    assert not torch.equal(st.master[64:128], w0[64:128])


def test_bce_from_logits_formula():
    x = torch.tensor([[-3.0], [0.0], [2.5], [40.0]])
    z = torch.tensor([0.0, 1.0, 1.0, 0.0])
    want = torch.clamp(x[:, 0], min=0) - x[:, 0] * z + torch.log1p(torch.exp(-x[:, 0].abs()))
    got = L.BinaryCrossentropy(from_logits=True).per_example(x, z)
    assert torch.allclose(got, want)
    assert torch.isfinite(got).all()


def test_cce_accepts_integer_and_one_hot_labels():
    x = torch.randn(6, 10)
    y = torch.randint(0, 10, (6,))
    l = L.CategoricalCrossentropy(from_logits=True)
    a = l.per_example(x, y)
    b = l.per_example(x, torch.nn.functional.one_hot(y, 10).float())
    assert torch.allclose(a, b, atol=1e-6)


def test_accuracy_thresholds_the_raw_logit_like_keras():
    logits = torch.tensor([[0.3], [0.7], [-0.2], [2.0]])
    labels = torch.tensor([1.0, 1.0, 0.0, 0.0])
    m = M.get("accuracy", 1)
    m.update(logits, labels)
    # logit 0.3 < 0.5 -> predicted 0 although sigmoid(0.3) > 0.5: the Keras quirk
    assert m.result() == 0.5
    p = M.BinaryAccuracy(threshold_on="prob")
    p.update(logits, labels)
    assert p.result() == 0.75


def test_auroc_matches_sklearn():
    sk = pytest.importorskip("sklearn.metrics")
    rng = np.random.RandomState(0)
    y = rng.randint(0, 2, 200)
    s = rng.randn(200) + y
    s[::7] = s[1::7][:len(s[::7])]      # ties
    assert abs(M.roc_auc_score(y, s) - sk.roc_auc_score(y, s)) < 1e-12


def test_history_and_initial_epoch_semantics():
    m = build_transfer_model("mobilenet_v2", device="cpu")
    m.compile(optimizer=O.RMSprop(1e-3), loss="bce", metrics=["accuracy"])
    x = torch.rand(8, 50, 50, 3)
    y = (torch.rand(8) > 0.5).float()
    h = m.fit([(x, y)], epochs=3, validation_data=[(x, y)], verbose=0)
    assert h.epoch == [0, 1, 2]
    assert set(h.history) == {"loss", "accuracy", "val_loss", "val_accuracy"}
    h2 = m.fit([(x, y)], epochs=5, initial_epoch=h.epoch[-1], verbose=0)
    assert h2.epoch == [2, 3, 4]        # reference: epochs=20, initial_epoch=9 -> 11 epochs
    res = m.evaluate([(x, y)] * 30, steps=20)
    assert len(res) == 2


def test_recompile_gives_fresh_optimizer_state():
    m = build_transfer_model("mobilenet_v2", device="cpu")
    m.compile(optimizer=O.RMSprop(1e-3), loss="bce", metrics=["accuracy"])
    x = torch.rand(4, 50, 50, 3)
    y = torch.tensor([0., 1., 1., 0.])
    m.fit([(x, y)], epochs=1, verbose=0)
    assert m.optimizer.s1.abs().sum() > 0
    m.compile(optimizer=O.RMSprop(1e-4), loss="bce", metrics=["accuracy"])
    assert m.optimizer.s1.abs().sum() == 0 and m.optimizer.iterations == 0


def test_dataset_ops_match_tf_data_semantics():
    ds = Dataset.from_tensor_slices(list(range(10)))
    assert list(ds.take(3)) == [0, 1, 2]
    assert list(ds.skip(7)) == [7, 8, 9]
    assert list(ds.shard(3, 1)) == [1, 4, 7]
    assert list(ds.take(2).concatenate(ds.skip(8))) == [0, 1, 8, 9]
    assert list(ds.take(2).repeat(2)) == [0, 1, 0, 1]
    b = list(ds.batch(4))
    assert [len(x) for x in b] == [4, 4, 2]               # no drop_remainder
    sh = list(ds.shuffle(4, seed=1))
    assert sorted(sh) == list(range(10)) and sh != list(range(10))
    calls = []
    c = ds.map(lambda v: (calls.append(v), v * 2)[1]).cache()
    assert list(c) == list(c) and len(calls) == 10        # second pass served from the cache
    assert list(ds.map(lambda v: v + 1, num_parallel_calls=4)) == list(range(1, 11))


def test_prepare_for_training_batches_synthetic_idc():
    ds = synthetic.synthetic_idc(70, seed=0)
    batches = list(prepare_for_training(ds, 32, seed=0))
    assert [b[0].shape[0] for b in batches] == [32, 32, 6]
    x, y = batches[0]
    assert x.shape == (32, 50, 50, 3) and x.dtype == torch.float32 and 0 <= x.min() and x.max() <= 1
    assert set(y.tolist()) <= {0.0, 1.0}
    srt = synthetic.synthetic_idc(10, order="class_sorted", label_as_int=True)
    labs = [l for _, l in srt]
    assert labs == sorted(labs, reverse=True)             # all class-1 first: the non-IID listing


def test_keras_style_callbacks(tmp_path):
    """EarlyStopping (with restore_best_weights), LearningRateScheduler, ReduceLROnPlateau, CSVLogger"""
    import numpy as np
    import torch
    from idc_models_b200.models import applications as apps
    from idc_models_b200.train import engine as E

    def model(lr=0.05):
        m = apps.Sequential([apps.Conv2D(8, (3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                             apps.Flatten(), apps.Dense(1)], device="cpu", seed=1)
        m.compile(optimizer=apps_opt(lr), loss="bce", metrics=["accuracy"])
        return m
    from idc_models_b200.optim.optimizers import SGD
    apps_opt = lambda lr: SGD(lr)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(32, 10, 10, 3, generator=g)
    y = (x.mean(dim=(1, 2, 3)) > 0.5).float()
    data = [(x[i:i + 8], y[i:i + 8]) for i in range(0, 32, 8)]
    # schedule: lr halves every epoch; the optimizer sees it and history records it
    m = model()
    h = m.fit(data, epochs=3, verbose=0, callbacks=[E.LearningRateScheduler(lambda epoch, lr: 0.05 * 0.5 ** epoch),
                                                    E.CSVLogger(str(tmp_path / "log.csv"))])
    assert h.history["lr"] == [0.05, 0.025, 0.0125] and m.optimizer.learning_rate == 0.0125
    rows = (tmp_path / "log.csv").read_text().strip().splitlines()
    assert rows[0].split(",")[0] == "epoch" and "loss" in rows[0] and len(rows) == 4
    # early stopping on a metric that cannot improve (lr 0): stops after `patience` non-improving epochs
    m = model(0.0)
    es = E.EarlyStopping(monitor="loss", patience=2, restore_best_weights=True)
    w0 = [w.copy() for w in m.get_weights()]
    h = m.fit(data, epochs=10, verbose=0, callbacks=[es])
    assert len(h.history["loss"]) == 3 and es.stopped_epoch == 2
    assert all(np.array_equal(a, b) for a, b in zip(w0, m.get_weights()))
    # plateau: a monitored value that never improves drives the learning rate down by `factor` every `patience`
    # epochs, floored at min_lr
    m = model(1.0)

    class Flat(E.Callback):
        def on_epoch_end(self, epoch, logs=None):
            logs["flat"] = 1.0
    rl = E.ReduceLROnPlateau(monitor="flat", factor=0.5, patience=1, min_lr=0.2, min_delta=0.0)
    h = m.fit(data, epochs=5, verbose=0, callbacks=[Flat(), rl])
    assert h.history["lr"] == [1.0, 1.0, 0.5, 0.25, 0.2] and m.optimizer.learning_rate == 0.2
    with pytest.raises(ValueError):
        E.ReduceLROnPlateau(factor=1.0)
