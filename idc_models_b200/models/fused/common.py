"""Shared machinery of the fused sm_100a executors.

An executor owns, per batch size, a *plan*: every activation / gradient / statistics buffer is
allocated once and reused by every step, so device addresses are stable (a requirement for CUDA-graph
capture) and a training step performs no allocation. All compute is launched through
``ops.conv`` / ``ops.kernels`` (our kernels); PyTorch is only used for buffer allocation, memsets and
host<->device copies.
"""
from __future__ import annotations

import torch

from ...ops import conv as OC
from ...ops import kernels as K
from ...ops import _lib

ACT_NONE, ACT_RELU, ACT_RELU6 = 0, 1, 2


class Scratch:
    """Bump allocator over one flat fp32 tensor that is zeroed with a single memset per step
    (per-channel statistics, reduction targets)."""

    def __init__(self, device):
        self.device = device
        self.items: list[tuple[int, int]] = []
        self.total = 0
        self.buf: torch.Tensor | None = None

    def reserve(self, n: int) -> int:
        off = self.total
        self.total += -(-n // 64) * 64
        return off

    def finalize(self) -> None:
        self.buf = torch.zeros(max(self.total, 64), dtype=torch.float32, device=self.device)

    def view(self, off: int, n: int) -> torch.Tensor:
        return self.buf[off:off + n]

    def zero(self) -> None:
        self.buf.zero_()


class InputStager:
    """Host -> device copy of a step's input on its own stream, double-buffered: the copy for step i+1 (issued by
    the host while step i is still running, which it is whenever the caller does not synchronise every step)
    overlaps the compute of step i instead of sitting in front of step i+1 on the compute stream. The graph reads
    its input from one fixed buffer, so the staged batch is moved there by a device-to-device copy on the compute
    stream (microseconds) once the H2D has landed. What ``tf.data``'s prefetch-to-device does for the reference's
    ``model.fit`` (/root/reference/dist_model_tf_mobile.py:63,136)."""

    def __init__(self, dst: torch.Tensor):
        self.dst = dst
        self.bufs = [torch.empty_like(dst) for _ in range(2)]
        self.stream = torch.cuda.Stream(device=dst.device)
        self.ready = [torch.cuda.Event() for _ in range(2)]     # H2D into buffer k has finished
        self.free = [torch.cuda.Event() for _ in range(2)]      # buffer k has been moved into dst
        self.k = 0

    def push(self, x: torch.Tensor) -> None:
        k = self.k
        self.k ^= 1
        cur = torch.cuda.current_stream(self.dst.device)
        self.stream.wait_event(self.free[k])        # (nothing to wait for until the event has been recorded once)
        with torch.cuda.stream(self.stream):
            self.bufs[k].copy_(x, non_blocking=True)
            self.ready[k].record(self.stream)
        cur.wait_event(self.ready[k])
        self.dst.copy_(self.bufs[k], non_blocking=True)
        self.free[k].record(cur)


class FusedExecutor:
    """Base class: input staging, loss/metric plumbing, plan cache, optional CUDA-graph replay."""

    # pinned host inputs are staged on a copy stream (InputStager); IDCB_ASYNC_STAGE=0 puts the copy back on the
    # compute stream
    async_stage = __import__("os").environ.get("IDCB_ASYNC_STAGE", "1") != "0"

    loss_kinds = {"bce": 0, "cce": 1}

    def __init__(self, net):
        if net.device.type != "cuda":
            raise RuntimeError("fused executors need a CUDA device")
        _lib.cuda_lib()  # fail loudly if the native library is missing
        self.net = net
        self.plans: dict = {}
        self.version = 0          # bumped whenever weights may have changed outside the optimizer
        self.first_trainable = None
        self.use_graphs = False
        self.capturing = False
        self._graphs: dict = {}
        # Set by a data-parallel strategy that overlaps the gradient all-reduce with the backward pass
        # (parallel/fused_comm.py:OverlapSchedule): begin() / ready(offset) / end() / key().
        self.grad_hook = None

    # ---- gradient-readiness protocol (buckets leave while backward continues) -----------
    def grad_ready_offsets(self) -> list:
        """Flat-store offsets, in the order the backward pass reaches them, such that when offset ``o`` is
        announced every parameter gradient at offsets >= ``o`` is final. Static per architecture (every
        rank derives the same bucket plan from it, also a rank whose local batch is empty)."""
        return []

    def grads_ready(self, offset: int) -> None:
        if self.grad_hook is not None and self._hook_live:
            self.grad_hook.ready(int(offset))

    def announcement_fires(self, offset: int) -> bool:
        """True if announcing ``offset`` now would launch a gradient bucket (lets an executor skip the side-stream
        join an announcement needs when the schedule would merge it into a later bucket anyway)."""
        h = self.grad_hook
        if h is None or not self._hook_live:
            return False
        fires = getattr(h, "fires", None)
        return bool(fires(int(offset))) if fires is not None else True

    _hook_live = False

    def _run_hooked(self, plan, training: bool, grad_scale: float) -> None:
        hook = self.grad_hook if training else None
        if hook is None:
            self.run(plan, training, grad_scale)
            return
        hook.scalars_src = plan.acc          # the step's metric slots ride on the last bucket's barrier
        hook.begin()
        self._hook_live = True
        try:
            self.run(plan, training, grad_scale)
        finally:
            self._hook_live = False
        hook.end()

    # ---- trainable bookkeeping -----------------------------------------------------------------
    def prepare(self) -> None:
        net = self.net
        idx = sorted(net.trainable_layer_indices() & {l.index for l in net.layers if l.spec.params})
        self.first_trainable = idx[0] if idx else len(net.layers)
        self.trainable = {l.index for l in net.layers if l.trainable}
        self.plans.clear()
        self._graphs.clear()

    def layer_trainable(self, name: str) -> bool:
        return self.net.get_layer(name).index in self.trainable

    def needs_grad_below(self, layer_index: int) -> bool:
        """True if some trainable layer sits at or before ``layer_index`` (so a gradient w.r.t. that
        layer's output must be propagated further down)."""
        return self.first_trainable <= layer_index

    # ---- weights -----------------------------------------------------------------------------------
    def w16(self, name: str) -> torch.Tensor:
        return self.net.store.storage_view(name, "shadow")

    def w32(self, name: str) -> torch.Tensor:
        return self.net.store.storage_view(name, "master")

    def gw(self, name: str) -> torch.Tensor:
        return self.net.store.storage_view(name, "grad")

    # ---- protocol -----------------------------------------------------------------------------------
    def _get_plan(self, batch: int):
        p = self.plans.get(batch)
        if p is None:
            p = self.build_plan(batch)
            self.plans[batch] = p
        return p

    def build_plan(self, batch: int):  # pragma: no cover - abstract
        raise NotImplementedError

    def run(self, plan, training: bool, grad_scale: float) -> None:  # pragma: no cover - abstract
        raise NotImplementedError

    def _stage(self, plan, x: torch.Tensor, y: torch.Tensor | None) -> None:
        if (self.async_stage and x.device.type == "cpu" and x.dtype == torch.float32 and x.is_contiguous()
                and x.numel() == plan.x32.numel() and x.is_pinned()):
            st = getattr(plan, "_stager", None)
            if st is None:
                st = plan._stager = InputStager(plan.x32)
            st.push(x.view(plan.x32.shape))
        else:
            plan.x32.copy_(x.to(torch.float32), non_blocking=True)
        if y is not None:
            plan.labels.copy_(y.reshape(-1).to(torch.float32), non_blocking=True)
        else:
            plan.labels.zero_()

    def _execute(self, plan, training: bool, grad_scale: float) -> None:
        # frozen-layer BatchNorm folds are baked into a captured graph, so the key carries the
        # external-write counter of the weights (and the overlap schedule's frozen scalars)
        hook = self.grad_hook if training else None
        hk = hook.key() if hook is not None else None
        key = (id(plan), training, grad_scale, self.net.store.ext_version, hk)
        if self.use_graphs:
            g = self._graphs.get(key)
            if g is None:
                # warm-up run (module loading, function attributes), then capture
                self._run_hooked(plan, training, grad_scale)
                if training and hook is None:
                    self.net.store.grad.zero_()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                n0 = _lib.launches()
                self.capturing = True
                try:
                    with torch.cuda.graph(g):
                        self._run_hooked(plan, training, grad_scale)
                finally:
                    self.capturing = False
                self._graphs[key] = (g, _lib.launches() - n0)
                if hook is not None:
                    # the warm-up run WAS this step (it reduced the gradients and applied the optimizer on
                    # every rank); the captured graph is first replayed by the next step
                    hook.consumed = True
                    return
                if training:
                    self.net.store.grad.zero_()
                g.replay()
                _lib.count_launch(self._graphs[key][1])
                return
            g[0].replay()
            _lib.count_launch(g[1])
        else:
            self._run_hooked(plan, training, grad_scale)
        if hook is not None:
            hook.consumed = True

    def train_batch(self, x, y, global_batch: int):
        b = x.shape[0]
        if b == 0:
            return torch.zeros(3, device=self.net.device), torch.zeros(0, self.net.n_outputs, device=self.net.device)
        plan = self._get_plan(b)
        self._stage(plan, x, y)
        self._execute(plan, True, 1.0 / float(global_batch))
        return plan.acc[:3], plan.logits

    def eval_batch(self, x, y):
        b = x.shape[0]
        if b == 0:
            return torch.zeros(3, device=self.net.device), torch.zeros(0, self.net.n_outputs, device=self.net.device)
        plan = self._get_plan(b)
        self._stage(plan, x, y)
        self._execute(plan, False, 0.0)
        return plan.acc[:3].clone(), plan.logits.clone()

    def predict(self, x):
        _, logits = self.eval_batch(x, None)
        return logits


class BNState:
    """Per-BatchNorm-layer device state used by the fused pipelines: scale/shift (the prologue
    constants), batch mean / rstd, and the sums the backward kernels fill."""

    def __init__(self, ex: FusedExecutor, name: str, channels: int, scratch: Scratch, need_bwd: bool):
        dev = ex.net.device
        self.ex = ex
        self._folded_at = None
        self.name = name
        self.c = channels
        sp = ex.net.get_layer(name).spec
        self.eps = sp.attrs["eps"]
        self.momentum = sp.attrs["momentum"]
        self.trainable = ex.layer_trainable(name)
        self.scale = torch.empty(channels, dtype=torch.float32, device=dev)
        self.shift = torch.empty(channels, dtype=torch.float32, device=dev)
        self.gamma = ex.w32(name + "/gamma")
        self.beta = ex.w32(name + "/beta")
        self.mmean = ex.w32(name + "/moving_mean")
        self.mvar = ex.w32(name + "/moving_variance")
        self.mean = torch.zeros(channels, dtype=torch.float32, device=dev)
        self.rstd = torch.zeros(channels, dtype=torch.float32, device=dev)
        self.o_sa = scratch.reserve(channels) if need_bwd else None
        self.o_sb = scratch.reserve(channels) if need_bwd else None
        self.scratch = scratch
        if need_bwd:
            self.A = torch.empty(channels, dtype=torch.float32, device=dev)
            self.B = torch.empty(channels, dtype=torch.float32, device=dev)
            self.C0 = torch.empty(channels, dtype=torch.float32, device=dev)
        self.dgamma = ex.gw(name + "/gamma") if self.trainable else None
        self.dbeta = ex.gw(name + "/beta") if self.trainable else None

    @property
    def sa(self):
        return self.scratch.view(self.o_sa, self.c)

    @property
    def sb(self):
        return self.scratch.view(self.o_sb, self.c)

    def fold(self) -> None:
        """inference mode: scale/shift from the moving statistics. Cached: a frozen layer only changes
        when weights are written from outside the optimizer, a trainable one after every update."""
        st = self.ex.net.store
        key = st.version if self.trainable else st.ext_version
        if self._folded_at == ("fold", key) and not (self.ex.capturing and self.trainable):
            return
        K.bn_fold(self.gamma, self.beta, self.mmean, self.mvar, self.eps, self.scale, self.shift)
        self._folded_at = ("fold", key)

    def finalize(self, sum_, sq, count: float) -> None:
        """training mode: scale/shift from batch sums; updates the moving statistics"""
        K.bn_finalize(sum_, sq, count, self.gamma, self.beta, self.eps, self.scale, self.shift,
                      self.mean, self.rstd, self.mmean, self.mvar, self.momentum)
        self._folded_at = None
