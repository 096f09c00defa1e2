"""Keras-flavoured optimizers over the flat parameter store.

An optimizer object only carries hyper-parameters (like a Keras optimizer before ``compile``); its
state is created by :meth:`Optimizer.bind` when a model is (re)compiled — recompiling therefore gives
FRESH state, which is what the reference's phase-2 ``model.compile(...)`` does
(/root/reference/dist_model_tf_mobile.py:152-154) and what TFF does per client per round
(fed_model.py:208).

``apply`` runs one fused launch per contiguous trainable range (``csrc/optim.cu``) on CUDA and the same
formulas in PyTorch on CPU.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from ..ops import _lib

KIND = {"rmsprop": 0, "sgd": 1, "adam": 2}


class Optimizer:
    kind = "sgd"

    def __init__(self, learning_rate: float = 0.01, lr: float | None = None, **kw):
        self.learning_rate = float(lr if lr is not None else learning_rate)
        self.store = None
        self.ranges: list[tuple[int, int]] = []
        self.s1 = self.s2 = None
        self.iterations = 0

    # hyper-parameters p1, p2, eps as consumed by the kernel
    def _params(self) -> tuple[float, float, float]:
        raise NotImplementedError

    def _n_state(self) -> int:
        return 1

    def _step_lr(self) -> float:
        return self.learning_rate

    def bind(self, store, ranges) -> "Optimizer":
        self.store = store
        self.ranges = list(ranges)
        n = store.master.numel()
        ns = self._n_state()
        self.s1 = torch.zeros(n, dtype=torch.float32, device=store.master.device) if ns >= 1 else None
        self.s2 = torch.zeros(n, dtype=torch.float32, device=store.master.device) if ns >= 2 else None
        self.iterations = 0
        return self

    def clone(self) -> "Optimizer":
        import copy
        o = copy.copy(self)
        o.store = None
        o.s1 = o.s2 = None
        o.ranges = []
        o.iterations = 0
        return o

    def apply(self, grad_scale: float = 1.0, zero_grad: bool = True, ranges=None) -> None:
        """w <- update(w, grad_scale * g) over the bound trainable ranges."""
        st = self.store
        self.iterations += 1
        if hasattr(st, "touch"):
            st.touch()
        lr = self._step_lr()
        p1, p2, eps = self._params()
        for b, e in (ranges if ranges is not None else self.ranges):
            if e <= b:
                continue
            if st.master.is_cuda:
                lib = _lib.cuda_lib()
                f4 = 4
                sh = st.shadow
                _lib.check(lib.idcb_optim_step(
                    KIND[self.kind], C.c_void_p(st.master.data_ptr() + b * f4),
                    C.c_void_p(st.grad.data_ptr() + b * f4),
                    C.c_void_p(self.s1.data_ptr() + b * f4 if self.s1 is not None else 0),
                    C.c_void_p(self.s2.data_ptr() + b * f4 if self.s2 is not None else 0),
                    C.c_void_p(sh.data_ptr() + b * 2 if sh is not None else 0),
                    C.c_longlong(e - b), C.c_float(lr), C.c_float(p1), C.c_float(p2), C.c_float(eps),
                    C.c_float(grad_scale), C.c_int(1 if zero_grad else 0), _lib.stream()))
                _lib.count_launch()
            else:
                self._apply_torch(b, e, lr, p1, p2, eps, grad_scale, zero_grad)

    def _apply_torch(self, b, e, lr, p1, p2, eps, grad_scale, zero_grad):
        st = self.store
        w = st.master[b:e]
        g = st.grad[b:e] * grad_scale
        if self.kind == "rmsprop":
            s1 = self.s1[b:e]
            s1.mul_(p1).addcmul_(g, g, value=1 - p1)
            w.sub_(lr * g / (s1.sqrt() + eps))
        elif self.kind == "sgd":
            s1 = self.s1[b:e]
            s1.mul_(p1).sub_(lr * g)
            w.add_(s1)
        else:
            s1, s2 = self.s1[b:e], self.s2[b:e]
            s1.mul_(p1).add_(g, alpha=1 - p1)
            s2.mul_(p2).addcmul_(g, g, value=1 - p2)
            w.sub_(lr * s1 / (s2.sqrt() + eps))
        if zero_grad:
            st.grad[b:e].zero_()


class RMSprop(Optimizer):
    """tf.keras.optimizers.RMSprop defaults: rho=0.9, momentum=0, epsilon=1e-7, centered=False."""
    kind = "rmsprop"

    def __init__(self, learning_rate=0.001, rho=0.9, epsilon=1e-7, lr=None, **kw):
        super().__init__(learning_rate, lr)
        self.rho, self.epsilon = float(rho), float(epsilon)

    def _params(self):
        return self.rho, 0.0, self.epsilon


class SGD(Optimizer):
    kind = "sgd"

    def __init__(self, learning_rate=0.01, momentum=0.0, lr=None, **kw):
        super().__init__(learning_rate, lr)
        self.momentum = float(momentum)

    def _params(self):
        return self.momentum, 0.0, 0.0


class Adam(Optimizer):
    """tf.keras.optimizers.Adam defaults: beta_1=0.9, beta_2=0.999, epsilon=1e-7."""
    kind = "adam"

    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, lr=None, **kw):
        super().__init__(learning_rate, lr)
        self.beta_1, self.beta_2, self.epsilon = float(beta_1), float(beta_2), float(epsilon)

    def _n_state(self):
        return 2

    def _step_lr(self):
        t = self.iterations
        return self.learning_rate * math.sqrt(1 - self.beta_2 ** t) / (1 - self.beta_1 ** t)

    def _params(self):
        return self.beta_1, self.beta_2, self.epsilon


def get(opt) -> Optimizer:
    if isinstance(opt, Optimizer):
        return opt
    table = {"rmsprop": RMSprop, "sgd": SGD, "adam": Adam}
    if isinstance(opt, str) and opt.lower() in table:
        return table[opt.lower()]()
    raise ValueError(f"unknown optimizer {opt!r}")
