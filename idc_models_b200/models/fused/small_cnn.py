"""Fused pipeline for the 10x10 secure-FL CNN (/root/reference/secure_fed_model.py:84-98): the whole
network — conv, pool, dropout, two dense layers, BCE loss, accuracy counters and the complete backward —
is ONE kernel launch per step (csrc/small_cnn.cu). fp32 end to end. SURVEY.md 2.E K14, 8.14."""
from __future__ import annotations

import ctypes as C

import torch

from ...ops import _lib
from .common import FusedExecutor


class _Plan:
    pass


class SmallCnnExecutor(FusedExecutor):
    def prepare(self) -> None:
        super().prepare()
        net = self.net
        kinds = [l.spec.kind for l in net.layers]
        if kinds != ["input", "conv", "maxpool", "dropout", "flatten", "dense", "dropout", "dense"] or \
                net.input_shape != (10, 10, 3) or net.n_outputs != 1:
            raise NotImplementedError("SmallCnnExecutor implements exactly the reference's create_model() network")
        if len(self.trainable) and len(self.trainable) != len(net.layers):
            if any(l.spec.params and l.index not in self.trainable for l in net.layers):
                raise NotImplementedError("the fused small-CNN kernel trains all six tensors or none")
        st = net.store
        names = [p.name for l in net.layers for p in l.spec.params]
        self.offsets = (C.c_int * 6)(*[st.slots[n].offset for n in names])
        self.cstride = st.slots[names[0]].storage_shape[-1]
        self.rates = [l.spec.attrs["rate"] for l in net.layers if l.spec.kind == "dropout"]
        self.step = 0

    def build_plan(self, batch: int):
        dev = self.net.device
        p = _Plan()
        p.x32 = torch.zeros(batch, 10, 10, 3, dtype=torch.float32, device=dev)
        p.labels = torch.zeros(batch, dtype=torch.float32, device=dev)
        p.acc = torch.zeros(4, dtype=torch.float32, device=dev)
        p.logits = torch.zeros(batch, 1, dtype=torch.float32, device=dev)
        p.batch = batch
        return p

    def run(self, p, training: bool, grad_scale: float) -> None:
        net = self.net
        st = net.store
        p.acc.zero_()
        self.step += 1
        lib = _lib.cuda_lib()
        seed = int(getattr(net, "dropout_seed", 0))
        _lib.check(lib.idcb_small_cnn_step(
            C.c_void_p(p.x32.data_ptr()), C.c_void_p(p.labels.data_ptr()), C.c_void_p(st.master.data_ptr()),
            C.c_void_p(st.grad.data_ptr()), C.c_void_p(p.logits.data_ptr()), C.c_void_p(p.acc.data_ptr()),
            C.c_int(1), C.c_int(p.batch), C.c_int(st.master.numel()), self.offsets, C.c_int(self.cstride),
            C.c_int(1 if training else 0), C.c_float(grad_scale), C.c_float(self.rates[0]), C.c_float(self.rates[1]),
            C.c_ulonglong(seed), C.c_uint(self.step), C.c_float(net.acc_threshold), C.c_int(1 if net.acc_on_prob else 0),
            _lib.stream()))
        _lib.count_launch()
