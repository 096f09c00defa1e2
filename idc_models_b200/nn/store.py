"""Flat parameter storage.

All trainable tensors of a network live in ONE fp32 "master" buffer (Keras layer order), with a
same-shaped fp32 gradient buffer and, on CUDA, a bf16 "shadow" copy that the compute kernels read
(the fused optimizer kernel refreshes the shadow in the same pass that updates the master). BatchNorm
moving statistics live in a second flat buffer. Every parameter is a view into these buffers, so

* the optimizer is one launch over a contiguous range instead of one per variable (the reference's
  Keras/TF stack launches one RMSprop kernel per variable: SURVEY.md 2.E K13),
* the data-parallel gradient all-reduce is one packed buffer, which is what
  ``MirroredStrategy``'s ``NcclAllReduce(num_packs=1)`` builds by copying,
* freezing ``layers[:fine_tune_at]`` leaves the trainable set as a contiguous suffix.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

ALIGN = 64  # elements; keeps every view 256-byte (fp32) / 128-byte (bf16) aligned for TMA / v4 access


@dataclass
class Slot:
    name: str
    kind: str
    shape: tuple
    offset: int      # element offset in its flat buffer
    numel: int
    layer: int       # index of the owning layer
    trainable_kind: bool   # False for BN moving statistics (lives in the stats buffer)
    fan_in: int = 0
    fan_out: int = 0
    storage_shape: tuple = ()   # conv kernels with Cin % 8 != 0 (RGB stems) are stored with Cin padded
                                # to 8 so the tensor-core conv path can read them; the pad lanes are
                                # zero and stay zero (their gradients are exactly zero)


def storage_shape_of(kind: str, shape: tuple) -> tuple:
    if kind == "conv_kernel" and shape[-1] % 8 != 0:
        return (*shape[:-1], -(-shape[-1] // 8) * 8)
    return tuple(shape)


def _numel(shape) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


class ParamStore:
    def __init__(self, layers, device="cpu", seed: int = 0):
        self.device = torch.device(device)
        self.slots: dict[str, Slot] = {}
        self.order: list[str] = []
        off_w = off_s = 0
        for li, l in enumerate(layers):
            for p in l.params:
                sshape = storage_shape_of(p.kind, tuple(p.shape))
                n = _numel(sshape)
                if p.trainable:
                    s = Slot(p.name, p.kind, tuple(p.shape), off_w, n, li, True, p.fan_in, p.fan_out,
                             sshape)
                    off_w += -(-n // ALIGN) * ALIGN
                else:
                    s = Slot(p.name, p.kind, tuple(p.shape), off_s, n, li, False, 0, 0, sshape)
                    off_s += -(-n // ALIGN) * ALIGN
                if p.name in self.slots:
                    raise ValueError(f"duplicate weight name {p.name}")
                self.slots[p.name] = s
                self.order.append(p.name)
        self.n_master = off_w
        self.n_stats = off_s
        self.master = torch.zeros(max(off_w, 1), dtype=torch.float32, device=self.device)
        self.grad = torch.zeros(max(off_w, 1), dtype=torch.float32, device=self.device)
        self.stats = torch.zeros(max(off_s, 1), dtype=torch.float32, device=self.device)
        self.shadow = (torch.zeros(max(off_w, 1), dtype=torch.bfloat16, device=self.device)
                       if self.device.type == "cuda" else None)
        self.init_(seed)

    # ---- views ---------------------------------------------------------------------------
    def _buf(self, s: Slot):
        return self.master if s.trainable_kind else self.stats

    @staticmethod
    def _logical(t: torch.Tensor, s: Slot) -> torch.Tensor:
        t = t[s.offset:s.offset + s.numel].view(s.storage_shape)
        return t if s.storage_shape == s.shape else t[..., :s.shape[-1]]

    def view(self, name: str) -> torch.Tensor:
        """Logical-shape view of a weight (a strided slice when the storage is channel-padded)."""
        s = self.slots[name]
        return self._logical(self._buf(s), s)

    def grad_view(self, name: str) -> torch.Tensor:
        s = self.slots[name]
        assert s.trainable_kind
        return self._logical(self.grad, s)

    def storage_view(self, name: str, which: str = "shadow") -> torch.Tensor:
        """Contiguous storage-shape view (what the kernels read): 'shadow' (bf16), 'master', 'grad'."""
        s = self.slots[name]
        buf = {"shadow": self.shadow, "master": self._buf(s), "grad": self.grad}[which]
        return buf[s.offset:s.offset + s.numel].view(s.storage_shape)

    def shadow_view(self, name: str) -> torch.Tensor:
        return self.storage_view(name, "shadow")

    # Change counters let executors cache values derived from weights (folded BatchNorm scale/shift):
    # ``version`` moves on every update, ``ext_version`` only when weights are written from outside the
    # optimizer (set_weights / load / broadcast / federated averaging), i.e. when frozen layers may change.
    version = 0
    ext_version = 0

    def touch(self, external: bool = False) -> None:
        self.version += 1
        if external:
            self.ext_version += 1

    def refresh_shadow(self) -> None:
        """Call after writing the master weights from outside the optimizer."""
        self.touch(external=True)
        if self.shadow is not None:
            self.shadow.copy_(self.master)

    # ---- init (Keras defaults: glorot_uniform kernels, zero bias, BN 1/0/0/1) -------------
    def init_(self, seed: int = 0) -> None:
        g = torch.Generator(device="cpu")
        g.manual_seed(seed)
        for name in self.order:
            s = self.slots[name]
            v = self.view(name)
            if s.kind in ("conv_kernel", "dw_kernel", "dense_kernel"):
                if s.kind == "dw_kernel":  # Keras DepthwiseConv2D: fan_in = kh*kw*C, fan_out = kh*kw*C*mult... uses full shape
                    kh, kw, c = s.shape
                    fan_in, fan_out = kh * kw * c, kh * kw
                else:
                    fan_in, fan_out = s.fan_in, s.fan_out
                lim = math.sqrt(6.0 / (fan_in + fan_out))
                t = (torch.rand(s.shape, generator=g) * 2 - 1) * lim
                v.copy_(t)
            elif s.kind in ("gamma", "moving_var"):
                v.fill_(1.0)
            else:
                v.zero_()
        self.refresh_shadow()

    # ---- trainable ranges ------------------------------------------------------------------
    def ranges(self, trainable_layers) -> list[tuple[int, int]]:
        """Maximal contiguous [begin, end) element ranges of the master buffer owned by layers whose
        index is in ``trainable_layers``. For ``layers[:fine_tune_at]`` freezing this is one range."""
        out: list[list[int]] = []
        for name in self.order:
            s = self.slots[name]
            if not s.trainable_kind or s.layer not in trainable_layers:
                continue
            b, e = s.offset, s.offset + -(-s.numel // ALIGN) * ALIGN
            if out and out[-1][1] == b:
                out[-1][1] = e
            else:
                out.append([b, e])
        return [(b, e) for b, e in out]

    def to(self, device) -> "ParamStore":
        device = torch.device(device)
        if device == self.device:
            return self
        self.device = device
        self.master = self.master.to(device)
        self.grad = self.grad.to(device)
        self.stats = self.stats.to(device)
        self.shadow = self.master.to(torch.bfloat16) if device.type == "cuda" else None
        return self
