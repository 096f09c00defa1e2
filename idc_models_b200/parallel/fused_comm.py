"""The fused data-parallel step: gradient all-reduce + loss-unscale + optimizer + bf16 cast + all-gather
in ONE kernel over NVLink peer / multicast memory (csrc/comm.cu:allreduce_optim_kernel).

Replaces what ``tf.distribute.MirroredStrategy`` does inside ``optimizer.apply_gradients`` — pack all
gradients, NCCL ring all-reduce, then one optimizer kernel per variable
(/root/reference/dist_model_tf_mobile.py:115-132). The NCCL route
stays available as ``MirroredStrategy(comm='nccl')`` and is the baseline this path is measured against.

On construction the model's flat master / gradient / bf16-shadow buffers are moved into symmetric
memory (every parameter is a view into them, so nothing else changes). Per step each rank:
reduces its 1/N shard of the gradient bucket straight out of every peer's memory (or through the
switch with ``multimem.ld_reduce``), applies the optimizer to that shard (optimizer state is therefore
sharded N ways for free), and writes the new fp32 weights and their bf16 cast into every peer (or
``multimem.st``). Replicas stay bit-identical because each value is computed once, by its owner.
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import _lib
from ..optim.optimizers import KIND
from .symm import SymmAllocator, ptr_array

vp = C.c_void_p

# launches of csrc/comm.cu:allreduce_optim_kernel issued (or captured) by this process; bench.py reports the
# per-step count next to its replica checksum
ALLREDUCE_LAUNCHES = 0


def _count_allreduce() -> None:
    global ALLREDUCE_LAUNCHES
    ALLREDUCE_LAUNCHES += 1
    _lib.count_launch()


class FusedComm:
    def __init__(self, world, store, use_multicast: bool = True):
        if world.device.type != "cuda":
            raise RuntimeError("the fused communication path needs CUDA devices (use comm='gloo' on CPU)")
        self.world = world
        self.store = store
        self.alloc = SymmAllocator(world, use_multicast)
        n = store.master.numel()
        self.master = self.alloc.alloc(n, torch.float32)
        self.grad = self.alloc.alloc(n, torch.float32)
        self.shadow = self.alloc.alloc(n, torch.bfloat16)
        self.master.tensor[:n].copy_(store.master)
        self.grad.tensor[:n].copy_(store.grad)
        self.shadow.tensor[:n].copy_(store.master.to(torch.bfloat16))
        # re-seat the store on the symmetric buffers (all parameter views are created from these)
        store.master = self.master.tensor[:n]
        store.grad = self.grad.tensor[:n]
        store.shadow = self.shadow.tensor[:n]
        store.touch(external=True)
        # BatchNorm moving statistics (averaged over replicas on read): same treatment, padded to 4 floats
        self.n_stats = int(store.stats.numel()) if getattr(store, "n_stats", 0) else 0
        if self.n_stats:
            self.stats = self.alloc.alloc(-(-self.n_stats // 4) * 4, torch.float32)
            self.stats.tensor[:self.n_stats].copy_(store.stats)
            store.stats = self.stats.tensor[:self.n_stats]
        self.scal = self.alloc.alloc(4, torch.float32)
        self.scal_out = torch.zeros(4, dtype=torch.float32, device=world.device)
        self.red = self.alloc.alloc(64, torch.float32)       # reduce_scalars staging (up to 64 values)
        torch.cuda.synchronize()
        world.barrier()

    @property
    def uses_multicast(self) -> bool:
        return bool(self.master.mc_ptr and self.grad.mc_ptr and self.shadow.mc_ptr)

    def allreduce_apply(self, optimizer, grad_scale: float = 1.0, owner: int = -1, scalars=None):
        """One launch per contiguous trainable range (normally one). ``owner`` >= 0 selects the central-
        storage variant (that rank alone reduces, updates and broadcasts). ``scalars`` (a 4-float device
        tensor) is summed over ranks on the kernel's barrier; the sum is returned."""
        want_scalars = scalars is not None
        if want_scalars:
            k = min(4, scalars.numel())
            self.scal.tensor[:k].copy_(scalars.reshape(-1)[:k])
        optimizer.iterations += 1
        self.store.touch()
        lr = optimizer._step_lr()
        p1, p2, eps = optimizer._params()
        lib = _lib.cuda_lib()
        cw = self.alloc.comm_world()
        mc = self.uses_multicast
        for b, e in optimizer.ranges:
            if e <= b:
                continue
            s1 = optimizer.s1
            s2 = optimizer.s2
            _lib.check(lib.idcb_allreduce_optim(
                C.byref(cw), self.grad.ptrs(b), self.master.ptrs(b), self.shadow.ptrs(b),
                self.grad.mc(b) if mc else vp(0), self.master.mc(b) if mc else vp(0),
                self.shadow.mc(b) if mc else vp(0),
                vp(s1.data_ptr() + 4 * b if s1 is not None else 0),
                vp(s2.data_ptr() + 4 * b if s2 is not None else 0),
                C.c_longlong(e - b), C.c_int(KIND[optimizer.kind]), C.c_float(lr), C.c_float(p1),
                C.c_float(p2), C.c_float(eps), C.c_float(grad_scale), C.c_int(owner),
                self.scal.ptrs() if scalars is not None else None,
                vp(self.scal_out.data_ptr()) if scalars is not None else vp(0), _lib.stream()))
            _count_allreduce()
            scalars = None      # reduce the metric slots once, with the first range
        return self.scal_out if want_scalars else None

    def reduce_scalars(self, t: torch.Tensor) -> torch.Tensor:
        """Sum of a handful of device scalars over ranks through the owner-computes kernel: no
        library collective. Values travel as fp32."""
        n = t.numel()
        if n > self.red.tensor.numel():
            raise ValueError("reduce_scalars: at most 64 values")
        n4 = -(-n // 4) * 4
        self.red.tensor[:n4].zero_()
        self.red.tensor[:n].copy_(t.reshape(-1).to(torch.float32))
        lib = _lib.cuda_lib()
        cw = self.alloc.comm_world()
        _lib.check(lib.idcb_stat_reduce(C.byref(cw), self.red.ptrs(), self.red.mc() if self.red.mc_ptr else vp(0),
                                        C.c_longlong(n4), C.c_float(1.0), _lib.stream()))
        _lib.count_launch()
        return self.red.tensor[:n].to(t.dtype).reshape(t.shape).clone()

    def stat_mean(self) -> None:
        """store.stats <- mean over ranks (csrc/comm.cu:stat_mean_kernel); every rank ends with the same values"""
        if not self.n_stats:
            return
        lib = _lib.cuda_lib()
        cw = self.alloc.comm_world()
        n4 = -(-self.n_stats // 4) * 4
        _lib.check(lib.idcb_stat_mean(C.byref(cw), self.stats.ptrs(), self.stats.mc() if self.stats.mc_ptr else vp(0),
                                      C.c_longlong(n4), _lib.stream()))
        _lib.count_launch()

    def allreduce_range(self, optimizer, b: int, e: int, grad_scale: float = 1.0, scalars=None) -> None:
        """One launch over the flat range [b, e) with the optimizer's CURRENT scalars; no step bookkeeping
        (used by :class:`OverlapSchedule`, which may run inside a captured graph). ``scalars``: a device
        tensor of >= 3 floats (or the string "zero") whose sum over ranks lands in ``self.scal_out``."""
        if e <= b:
            return
        if scalars is not None:
            if isinstance(scalars, str):
                self.scal.tensor[:4].zero_()
            else:
                k = min(4, scalars.numel())
                self.scal.tensor[:k].copy_(scalars.reshape(-1)[:k])
        lr = optimizer._step_lr()
        p1, p2, eps = optimizer._params()
        lib = _lib.cuda_lib()
        cw = self.alloc.comm_world()
        mc = self.uses_multicast
        s1, s2 = optimizer.s1, optimizer.s2
        _lib.check(lib.idcb_allreduce_optim(
            C.byref(cw), self.grad.ptrs(b), self.master.ptrs(b), self.shadow.ptrs(b),
            self.grad.mc(b) if mc else vp(0), self.master.mc(b) if mc else vp(0),
            self.shadow.mc(b) if mc else vp(0),
            vp(s1.data_ptr() + 4 * b if s1 is not None else 0),
            vp(s2.data_ptr() + 4 * b if s2 is not None else 0),
            C.c_longlong(e - b), C.c_int(KIND[optimizer.kind]), C.c_float(lr), C.c_float(p1),
            C.c_float(p2), C.c_float(eps), C.c_float(grad_scale), C.c_int(-1),
            self.scal.ptrs() if scalars is not None else None,
            vp(self.scal_out.data_ptr()) if scalars is not None else vp(0), _lib.stream()))
        _count_allreduce()


def plan_buckets(ready_offsets, lo: int, hi: int, min_bucket: int) -> list:
    """Bucket plan for the trainable range [lo, hi): ``ready_offsets`` are announced top-down by the backward
    pass; announcements are merged until a bucket holds at least ``min_bucket`` elements (each bucket costs a
    launch and two cross-GPU barriers). Returns ``[(announce_offset, begin, end)]`` in launch order; the last
    entry is whatever is left when the backward pass is done. Pure function of static data: every rank computes
    the same plan."""
    plan = []
    top = hi
    for o in ready_offsets:
        o = max(lo, min(int(o) & ~3, top))
        if top - o >= min_bucket and o > lo:
            plan.append((o, o, top))
            top = o
    plan.append((lo, lo, top))
    return plan


class OverlapSchedule:
    """Bucketed launch of the fused all-reduce+optimizer kernel WHILE the backward pass runs (SURVEY 8.1,
    7.2 step 5). The executor announces "every gradient at flat offset >= o is final" as its backward walks
    down the network; each announcement that closes a planned bucket forks a side stream off the compute
    stream (an event edge, also inside a captured CUDA graph), launches the kernel for that bucket there, and
    the compute stream joins the side stream when the step ends. The bucket plan is derived from the
    executor's static offset list, so every rank — also one whose local batch is empty — launches the same
    kernels in the same order (the in-kernel barriers pair up launch by launch).

    Hyper-parameters are kernel arguments and are therefore frozen into a captured graph: the schedule is
    offered for optimizers whose per-step scalars do not change (SGD, RMSprop with a constant learning rate);
    ``key()`` makes the executor re-capture if they do."""

    def __init__(self, fused: FusedComm, optimizer, executor, grad_scale: float = 1.0,
                 min_bucket: int = 1 << 19, max_ctas: int = 0):
        self.fused, self.optimizer, self.grad_scale = fused, optimizer, float(grad_scale)
        if optimizer.kind == "adam":
            raise ValueError("overlap needs step-invariant optimizer scalars (Adam's bias correction changes per step)")
        if len(optimizer.ranges) != 1:
            raise ValueError("overlap needs one contiguous trainable range")
        self.lo, self.hi = optimizer.ranges[0]
        self.plan = plan_buckets(executor.grad_ready_offsets(), self.lo, self.hi, min_bucket)
        import os
        self.overlap_ctas = int(os.environ.get("IDCB_OVERLAP_CTAS", "48"))
        self.side = torch.cuda.Stream(device=fused.world.device, priority=-1)
        self._next = 0
        self._forked = False
        self.consumed = False        # set by the executor once a training step has launched (or replayed) the buckets
        # the step's (loss sum, correct, count) slots on this rank: summed over ranks by the LAST bucket's kernel
        # (set by the executor before begin(); None = this rank contributes zeros)
        self.scalars_src = None

    def key(self):
        o = self.optimizer
        return (o.kind, o._step_lr(), o._params(), self.grad_scale, len(self.plan))

    def begin(self) -> None:
        self._next = 0
        self._forked = False

    def _launch(self, b: int, e: int, last: bool = False) -> None:
        cur = torch.cuda.current_stream()
        self.side.wait_stream(cur)                      # bucket's gradients are complete on the compute stream
        lib = _lib.cuda_lib()
        # buckets that run next to the backward kernels use few CTAs; the last one has the GPU to itself
        lib.idcb_comm_set_max_ctas(C.c_int(0 if last else self.overlap_ctas))
        with torch.cuda.stream(self.side):
            self.fused.allreduce_range(self.optimizer, b, e, self.grad_scale,
                                       scalars=(self.scalars_src if self.scalars_src is not None else "zero") if last else None)
        lib.idcb_comm_set_max_ctas(C.c_int(0))
        self._forked = True

    def fires(self, offset: int) -> bool:
        return self._next < len(self.plan) - 1 and offset <= self.plan[self._next][0]

    def ready(self, offset: int) -> None:
        while self._next < len(self.plan) - 1 and offset <= self.plan[self._next][0]:
            _, b, e = self.plan[self._next]
            self._launch(b, e)
            self._next += 1

    def end(self) -> None:
        """launch what is left (at least the bottom bucket) and join the side stream"""
        while self._next < len(self.plan):
            _, b, e = self.plan[self._next]
            self._launch(b, e, last=self._next == len(self.plan) - 1)
            self._next += 1
        if self._forked:
            torch.cuda.current_stream().wait_stream(self.side)

    def run_all(self) -> None:
        """a rank with an empty local batch: same launches, nothing to overlap with"""
        self.scalars_src = None
        self.begin()
        self.end()


def allreduce_apply_emulated(emu, rank_bufs, opt_states, n, kind, lr, p1, p2, eps, grad_scale, owner=-1) -> None:
    """Single-process emulation used by the 1-GPU tests: ``rank_bufs`` = dict(grad/master/shadow ->
    list[SymmBuffer] per rank); runs the kernel once per emulated rank with barriers disabled."""
    lib = _lib.cuda_lib()
    for r in range(emu.size):
        cw = emu.comm_world(r)
        s1, s2 = opt_states[r]
        _lib.check(lib.idcb_allreduce_optim(
            C.byref(cw), rank_bufs["grad"][r].ptrs(), rank_bufs["master"][r].ptrs(),
            rank_bufs["shadow"][r].ptrs(), vp(0), vp(0), vp(0),
            vp(s1.data_ptr() if s1 is not None else 0), vp(s2.data_ptr() if s2 is not None else 0),
            C.c_longlong(n), C.c_int(KIND[kind]), C.c_float(lr), C.c_float(p1), C.c_float(p2),
            C.c_float(eps), C.c_float(grad_scale), C.c_int(owner), None, vp(0), _lib.stream()))
        _count_allreduce()
