"""Native input pipeline: ``cache() -> shuffle(buffer) -> batch() -> prefetch()`` of the reference's
``prepare_for_training`` (/root/reference/dist_model_tf_mobile.py:47-65; SURVEY.md 2.A A4, 2.C "tf.data C++
runtime") executed by C++ worker threads (csrc/host/loader.cpp) into a ring of pinned host slots, plus an
asynchronous host->device stage on its own CUDA stream so the copy of batch k+1 overlaps the training step on
batch k.

    loader = NativeLoader.from_dataset(labeled_ds, batch_size=32, shuffle_buffer_size=1000, device="cuda")
    for x, y in loader:            # one epoch; a new shuffle order every epoch (tf.data reshuffles)
        model.train_on_batch(x, y)

The Python ``Dataset`` pipeline (data/dataset.py) stays the general path (arbitrary ``map`` functions, PNG
decode); this class takes over once the decoded samples are cached, which is where the reference spends all
but its first epoch.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..ops import _lib


def _bind(lib):
    if getattr(lib, "_loader_bound", False):
        return
    vp, ll, ci = C.c_void_p, C.c_longlong, C.c_int
    lib.idcbh_loader_create.restype = vp
    lib.idcbh_loader_create.argtypes = [vp, vp, ll, ll, ci, ci, ci, ci, C.POINTER(vp), C.POINTER(vp), ci, C.c_ulonglong]
    lib.idcbh_loader_start_epoch.argtypes = [vp]
    lib.idcbh_loader_next.argtypes = [vp, C.POINTER(ci), C.POINTER(ci)]
    lib.idcbh_loader_release.argtypes = [vp, ci]
    lib.idcbh_loader_num_batches.argtypes = [vp]
    lib.idcbh_loader_num_batches.restype = ll
    lib.idcbh_loader_destroy.argtypes = [vp]
    lib.idcbh_loader_destroy.restype = None
    lib._loader_bound = True


class NativeLoader:
    def __init__(self, images, labels, batch_size: int, shuffle_buffer_size: int = 1000, drop_remainder: bool = False,
                 seed: int = 0, n_slots: int = 4, n_threads: int = 2, device=None):
        """``images`` [N, ...] float32 and ``labels`` [N] (numpy or CPU torch); both are kept alive here."""
        self.images = torch.as_tensor(np.ascontiguousarray(images) if isinstance(images, np.ndarray) else images)
        self.images = self.images.to(torch.float32).contiguous()
        self.labels = torch.as_tensor(np.ascontiguousarray(labels) if isinstance(labels, np.ndarray) else labels)
        self.labels = self.labels.to(torch.float32).reshape(-1).contiguous()
        if self.images.device.type != "cpu" or self.labels.shape[0] != self.images.shape[0]:
            raise ValueError("NativeLoader needs CPU tensors with one label per sample")
        self.n = int(self.images.shape[0])
        self.sample_shape = tuple(self.images.shape[1:])
        self.sample_floats = int(np.prod(self.sample_shape)) if self.sample_shape else 1
        self.batch_size, self.n_slots = int(batch_size), int(n_slots)
        self.device = torch.device(device) if device is not None else None
        pin = torch.cuda.is_available()
        self.slot_x = [torch.empty((batch_size,) + self.sample_shape, dtype=torch.float32, pin_memory=pin)
                       for _ in range(n_slots)]
        self.slot_y = [torch.empty(batch_size, dtype=torch.float32, pin_memory=pin) for _ in range(n_slots)]
        lib = _lib.host_lib()
        _bind(lib)
        self._lib = lib
        ax = (C.c_void_p * n_slots)(*[t.data_ptr() for t in self.slot_x])
        ay = (C.c_void_p * n_slots)(*[t.data_ptr() for t in self.slot_y])
        self._h = lib.idcbh_loader_create(self.images.data_ptr(), self.labels.data_ptr(), self.n, self.sample_floats,
                                          batch_size, int(shuffle_buffer_size), int(bool(drop_remainder)), n_slots,
                                          ax, ay, int(n_threads), int(seed) & 0xFFFFFFFFFFFFFFFF)
        if not self._h:
            raise RuntimeError(lib.idcbh_last_error().decode())
        self.drop_remainder = bool(drop_remainder)
        self._copy_stream = None
        if self.device is not None and self.device.type == "cuda":
            self._copy_stream = torch.cuda.Stream(device=self.device)

    # ---- construction from the tf.data-like pipeline --------------------------------------------------------
    @classmethod
    def from_dataset(cls, ds, batch_size: int, shuffle_buffer_size: int = 1000, **kw) -> "NativeLoader":
        """Materialise (= ``cache()``) the elements of an (image, label) dataset and serve them natively."""
        xs, ys = [], []
        for x, y in ds:
            xs.append(np.asarray(x, dtype=np.float32))
            ys.append(float(np.asarray(y).reshape(-1)[0]))
        if not xs:
            raise ValueError("empty dataset")
        return cls(np.stack(xs), np.asarray(ys, dtype=np.float32), batch_size, shuffle_buffer_size, **kw)

    def __len__(self) -> int:
        full, rem = divmod(self.n, self.batch_size)
        return full + (1 if rem and not self.drop_remainder else 0)

    # ---- iteration: one epoch --------------------------------------------------------------------------------
    def _host_batches(self):
        lib, h = self._lib, self._h
        if lib.idcbh_loader_start_epoch(h) != 0:
            raise RuntimeError("loader: start_epoch failed")
        slot, cnt = C.c_int(), C.c_int()
        while True:
            r = lib.idcbh_loader_next(h, C.byref(slot), C.byref(cnt))
            if r == 0:
                return
            if r < 0:
                raise RuntimeError("loader: next failed")
            yield slot.value, cnt.value

    def __iter__(self):
        lib, h = self._lib, self._h
        if self._copy_stream is None:
            # host consumer: hand out copies so the slot can be recycled immediately
            for s, c in self._host_batches():
                x, y = self.slot_x[s][:c].clone(), self.slot_y[s][:c].clone()
                lib.idcbh_loader_release(h, s)
                yield x, y
            return
        # device consumer: H2D of batch k+1 is in flight on the copy stream while batch k is being used
        dev, cs = self.device, self._copy_stream
        pending = None            # (x_dev, y_dev, event, slot)
        try:
            for s, c in self._host_batches():
                with torch.cuda.stream(cs):
                    xd = self.slot_x[s][:c].to(dev, non_blocking=True)
                    yd = self.slot_y[s][:c].to(dev, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(cs)
                prev, pending = pending, (xd, yd, ev, s)
                if prev is not None:
                    px, py, pev, ps = prev
                    pev.synchronize()                          # its copy has left the pinned slot
                    lib.idcbh_loader_release(h, ps)
                    torch.cuda.current_stream(dev).wait_event(pev)
                    px.record_stream(torch.cuda.current_stream(dev))
                    py.record_stream(torch.cuda.current_stream(dev))
                    yield px, py
            if pending is not None:
                px, py, pev, ps = pending
                pending = None
                pev.synchronize()
                lib.idcbh_loader_release(h, ps)
                torch.cuda.current_stream(dev).wait_event(pev)
                px.record_stream(torch.cuda.current_stream(dev))
                py.record_stream(torch.cuda.current_stream(dev))
                yield px, py
        finally:
            # the consumer stopped early (steps_per_epoch, evaluate(steps=...)): the slot whose copy was
            # already issued must go back to the ring or the producer threads stall on the next epoch
            if pending is not None and getattr(self, "_h", None):
                pending[2].synchronize()
                lib.idcbh_loader_release(h, pending[3])

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.idcbh_loader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
