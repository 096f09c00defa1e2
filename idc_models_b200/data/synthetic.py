"""Synthetic stand-ins for the datasets the reference reads (there is no network, and the IDC PNG
folders / tfds CIFAR-10 are not on the box): same shapes, dtypes and label conventions as
``<path>/data/IDC_regular_ps50_idx5`` / ``balanced_IDC_30k`` (dist_model_tf_mobile.py:105,
dist_model_tf_vgg.py:105, fed_model.py:159-164) and ``tfds.load('cifar10')``
(dist_model_tf_dense.py:120).

Images carry a weak class-dependent signal (IDC-positive patches are darker/more purple on average,
CIFAR classes get a class-specific low-frequency pattern) so that a few training steps move the
accuracy above chance in smoke tests. Generation is a counter-based hash (deterministic per index),
identical on every rank, so data-parallel replicas can slice the same global batch.
"""
from __future__ import annotations

import numpy as np
import torch

from .dataset import Dataset


def _patch(idx: int, label: int, shape, seed: int, n_classes: int = 2) -> np.ndarray:
    h, w, c = shape
    rng = np.random.default_rng((seed * 1000003 + idx) & 0x7FFFFFFF)
    img = rng.random((h, w, c), dtype=np.float32)
    if n_classes == 2:
        tint = np.array([0.55, 0.35, 0.60], dtype=np.float32) if label else np.array([0.85, 0.70, 0.80], dtype=np.float32)
        img = 0.55 * img + 0.45 * tint[:c]
    else:
        yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
        pat = 0.5 + 0.5 * np.sin((label + 1) * 0.37 * xx / w * 6.28 + label) * np.cos((label + 2) * 0.21 * yy / h * 6.28)
        img = 0.6 * img + 0.4 * pat[..., None]
    return img.astype(np.float32)


def synthetic_idc(n: int, image_shape=(50, 50, 3), seed: int = 0, order: str = "shuffled",
                  label_as_int: bool = False) -> Dataset:
    """Balanced binary IDC-like patches. ``order``: 'shuffled' (IID file listing) or 'class_sorted'
    (all class-1 examples before all class-0 examples — the non-IID listing of fed_model.py:161-164)."""
    labels = (np.arange(n) % 2).astype(np.int64)
    if order == "shuffled":
        labels = labels[np.random.default_rng(seed).permutation(n)]
    elif order == "class_sorted":
        labels = np.concatenate([np.ones(n - n // 2, np.int64), np.zeros(n // 2, np.int64)])
    else:
        raise ValueError(order)

    def gen():
        for i in range(n):
            lab = int(labels[i])
            yield _patch(i, lab, image_shape, seed), (lab if label_as_int else bool(lab))
    return Dataset(gen, n)


def synthetic_cifar10(n: int, seed: int = 0, as_uint8: bool = True) -> Dataset:
    """32x32x3 uint8 images with int labels, like tfds cifar10 as_supervised=True."""
    def gen():
        for i in range(n):
            lab = int((i * 7 + seed) % 10)
            img = _patch(i, lab, (32, 32, 3), seed, n_classes=10)
            yield ((img * 255).astype(np.uint8) if as_uint8 else img), lab
    return Dataset(gen, n)


def scale(image, label):
    """cast to float32 and divide by 255 (dist_model_tf_dense.py:114-117)."""
    return np.asarray(image, dtype=np.float32) / 255.0, label


def synthetic_batch(batch: int, image_shape=(50, 50, 3), n_classes: int = 2, seed: int = 0,
                    device="cpu", pin: bool = False):
    """One dense random batch (benchmarks): float32 images in [0,1] and labels."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand((batch, *image_shape), generator=g)
    y = torch.randint(0, n_classes, (batch,), generator=g).to(torch.float32)
    if pin and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    if str(device) != "cpu":
        x, y = x.to(device), y.to(device)
    return x, y


class DeviceSyntheticIDC:
    """An endless-epoch iterable of synthetic IDC batches generated ON the GPU by a Philox kernel
    (csrc/small_cnn.cu:synth_patches_kernel): no host pipeline, no H2D copy. Every rank that
    uses the same ``seed`` produces the same global batches, so data-parallel replicas can slice them."""

    def __init__(self, steps_per_epoch: int, batch_size: int, image_shape=(50, 50, 3), seed: int = 0, device="cuda"):
        self.steps, self.batch, self.shape, self.seed = steps_per_epoch, batch_size, tuple(image_shape), seed
        self.device = torch.device(device)
        self._epoch = 0

    def __len__(self):
        return self.steps

    def __iter__(self):
        import ctypes as C
        from ..ops import _lib
        lib = _lib.cuda_lib()
        h, w, c = self.shape
        base = self._epoch * self.steps
        self._epoch += 1
        for i in range(self.steps):
            x = torch.empty(self.batch, h, w, c, dtype=torch.float32, device=self.device)
            y = torch.empty(self.batch, dtype=torch.float32, device=self.device)
            _lib.check(lib.idcb_synth_patches(C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), C.c_longlong(self.batch),
                                              C.c_int(h), C.c_int(w), C.c_int(c), C.c_ulonglong(self.seed),
                                              C.c_uint(base + i), _lib.stream()))
            _lib.count_launch()
            yield x, y
