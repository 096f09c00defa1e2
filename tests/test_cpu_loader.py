"""The native (C++ threads, pinned slot ring) input pipeline: tf.data semantics of
cache -> shuffle(buffer) -> batch -> prefetch (/root/reference/dist_model_tf_mobile.py:47-65)."""
import numpy as np
import torch

from idc_models_b200.data.dataset import Dataset
from idc_models_b200.data.loader import NativeLoader


def _data(n=103, shape=(4, 4, 3)):
    x = np.arange(n, dtype=np.float32)[:, None, None, None] * np.ones(shape, np.float32)
    y = (np.arange(n) % 2).astype(np.float32)
    return x, y


def _epoch(loader):
    ids, labels, sizes = [], [], []
    for xb, yb in loader:
        assert xb.dtype == torch.float32 and xb.shape[1:] == (4, 4, 3)
        assert torch.all(xb == xb[:, :1, :1, :1])            # samples arrive intact
        ids += xb[:, 0, 0, 0].long().tolist()
        labels += yb.tolist()
        sizes.append(xb.shape[0])
    return ids, labels, sizes


def test_every_sample_once_per_epoch_and_partial_last_batch():
    x, y = _data()
    ld = NativeLoader(x, y, batch_size=32, shuffle_buffer_size=1000, seed=1, n_slots=3, n_threads=3)
    assert len(ld) == 4
    ids, labels, sizes = _epoch(ld)
    assert sizes == [32, 32, 32, 7]
    assert sorted(ids) == list(range(103))
    assert labels == [float(i % 2) for i in ids]              # labels travel with their samples
    ids2, _, _ = _epoch(ld)
    assert sorted(ids2) == list(range(103)) and ids2 != ids   # reshuffled every epoch
    ld.close()


def test_drop_remainder_and_determinism():
    x, y = _data()
    a = NativeLoader(x, y, 32, 1000, drop_remainder=True, seed=7)
    b = NativeLoader(x, y, 32, 1000, drop_remainder=True, seed=7)
    ia, _, sa = _epoch(a)
    ib, _, _ = _epoch(b)
    assert sa == [32, 32, 32] and len(a) == 3
    assert ia == ib                                            # same seed, same order
    c = NativeLoader(x, y, 32, 1000, drop_remainder=True, seed=8)
    assert _epoch(c)[0] != ia


def test_shuffle_buffer_semantics():
    """tf.data shuffle(buffer): the element emitted at position p was among the first p + buffer inputs;
    buffer 1 is the identity."""
    x, y = _data(200)
    ids, _, _ = _epoch(NativeLoader(x, y, 16, shuffle_buffer_size=1, seed=3))
    assert ids == list(range(200))
    B = 10
    ids, _, _ = _epoch(NativeLoader(x, y, 16, shuffle_buffer_size=B, seed=3))
    assert sorted(ids) == list(range(200)) and ids != list(range(200))
    assert all(v < p + B for p, v in enumerate(ids))


def test_abandoned_epoch_then_restart():
    x, y = _data(300)
    ld = NativeLoader(x, y, 8, 50, seed=5, n_slots=2, n_threads=4)
    it = iter(ld)
    for _ in range(3):
        next(it)
    del it                                                      # e.g. steps_per_epoch reached
    ids, _, sizes = _epoch(ld)
    assert sorted(ids) == list(range(300)) and len(sizes) == 38


def test_from_dataset_matches_python_pipeline_contents():
    x, y = _data(40)
    ds = Dataset.from_tensor_slices((x, y))
    ld = NativeLoader.from_dataset(ds, batch_size=16, shuffle_buffer_size=1000, seed=0)
    ids, labels, sizes = _epoch(ld)
    assert sizes == [16, 16, 8] and sorted(ids) == list(range(40))
    assert labels == [float(i % 2) for i in ids]


def test_loader_ring_protocol_under_thread_sanitizer(tmp_path):
    """tests/native/tsan_loader.cpp: 4 worker threads against a consumer that holds two slots at a time, six epochs
    (two abandoned half way), destruction with workers waiting and filling — compiled with -fsanitize=thread, so a
    data race in the slot protocol is a report and a non-zero exit (host side)."""
    import shutil
    import subprocess
    from pathlib import Path
    import pytest
    root = Path(__file__).resolve().parent.parent
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    exe = tmp_path / "tsan_loader"
    host = root / "csrc" / "host"
    cmd = [cxx, "-O1", "-g", "-std=c++17", "-fsanitize=thread", "-pthread", "-I", str(root / "include"), "-I", str(host),
           str(host / "loader.cpp"), str(host / "host_runtime.cpp"), str(root / "tests" / "native" / "tsan_loader.cpp"),
           "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and any(t in r.stderr.lower() for t in ("sanitize", "asan", "ubsan", "tsan")):
        pytest.skip("this g++ has no thread-sanitizer runtime")
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    if "FATAL: ThreadSanitizer" in r.stderr and "memory layout" in r.stderr:
        pytest.skip("thread sanitizer cannot map its shadow memory on this kernel")
    assert r.returncode == 0 and "loader ok" in r.stdout and "WARNING: ThreadSanitizer" not in r.stderr, \
        (r.stdout[-500:], r.stderr[-3000:])
