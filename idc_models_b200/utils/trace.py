"""Observability beyond the reference's ``Timer`` prints:

* :func:`nvtx_range` – NVTX ranges around epochs / steps / rounds (visible in Nsight tools; a no-op on CPU);
* :class:`JsonlLogger` – one JSON record per epoch / round / benchmark step (loss, metrics, patches/s, device
  milliseconds), the machine-readable companion of the raw ``history`` dict the reference prints;
* :class:`DeviceTimer` – CUDA-event timing of a region on the current stream (device time, not host wall clock).

Logging is off unless a path is given (``--log-jsonl`` flag of the scripts or ``IDCB_LOG_JSONL``).
"""
from __future__ import annotations

import contextlib
import json
import os
import time

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class JsonlLogger:
    def __init__(self, path: str | None = None, rank: int = 0):
        path = path or os.environ.get("IDCB_LOG_JSONL")
        self.path = path if (path and rank == 0) else None
        if self.path:
            os.makedirs(os.path.dirname(os.path.abspath(self.path)) or ".", exist_ok=True)

    def log(self, kind: str, **fields) -> None:
        if not self.path:
            return
        rec = {"t": round(time.time(), 3), "kind": kind}
        rec.update({k: (float(v) if isinstance(v, (torch.Tensor,)) else v) for k, v in fields.items()})
        with open(self.path, "a") as fh:
            fh.write(json.dumps(rec) + "\n")


class DeviceTimer:
    """``with DeviceTimer() as t: ...`` then ``t.ms`` (synchronises on exit; CPU: wall clock)."""

    def __enter__(self):
        self.cuda = torch.cuda.is_available()
        if self.cuda:
            self.s, self.e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.s.record()
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self.e.record()
            self.e.synchronize()
            self.ms = self.s.elapsed_time(self.e)
        else:
            self.ms = (time.perf_counter() - self.t0) * 1e3
        return False
