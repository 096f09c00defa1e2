"""``Timer`` — the reference's only instrumentation (/root/reference/dist_model_tf_dense.py:31-44 and its
copies): a context manager printing ``"{name} took {seconds} seconds"`` (host wall-clock). The printed
line is part of the scripts' observable output and is kept verbatim; ``device=True`` additionally
synchronises the GPU so the number covers the queued device work."""
from __future__ import annotations

import time


class Timer(object):
    def __init__(self, name: str, device: bool = True, printer=print):
        self.name = name
        self.device = device
        self.printer = printer
        self.elapsed = None

    def _sync(self):
        if self.device:
            try:
                import torch
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
            except Exception:  # noqa: BLE001
                pass

    def __enter__(self):
        self._sync()
        self.t = time.time()
        return self

    def __exit__(self, *args, **kwargs):
        self._sync()
        self.elapsed = time.time() - self.t
        self.printer("{} took {} seconds".format(self.name, self.elapsed))
