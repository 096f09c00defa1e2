"""Symmetric (peer-mapped) device buffers for the fused communication kernels.

A :class:`SymmBuffer` is the same-sized allocation on every rank, mapped into every process; the
kernels receive the table of per-rank base pointers (``peer_ptrs``) and, when the fabric offers NVLS,
the multicast alias (``mc_ptr``) on which ``multimem.ld_reduce`` / ``multimem.st`` operate.

Allocation and the cross-process handle exchange are delegated to
``torch.distributed._symmetric_memory`` (CUDA VMM + fabric/fd handles). All data movement is done by
our own kernels (csrc/comm.cu) — torch is used only as the allocator/rendezvous.

Two degenerate worlds share the interface so that everything above is testable:
* world == 1: a plain tensor, one pointer, no multicast;
* :class:`EmulatedWorld`: N "ranks" held by ONE process on one GPU (N plain tensors); the kernels are
  run once per emulated rank with the in-kernel barriers disabled — valid because ranks own disjoint
  shards (see include/idcb200.h, ``idcb_comm_world.use_barrier``).
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import _lib

vp = C.c_void_p
MAX_WORLD = 8


class CommWorld(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("use_barrier", C.c_int),
                ("signal_pads", vp * MAX_WORLD)]


def ptr_array(ptrs) -> C.Array:
    arr = (vp * MAX_WORLD)()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr


class SymmBuffer:
    """One symmetric allocation: ``tensor`` is this rank's copy."""

    def __init__(self, tensor: torch.Tensor, peer_ptrs: list, mc_ptr: int = 0, handle=None):
        self.tensor = tensor
        self.peer_ptrs = list(peer_ptrs)
        self.mc_ptr = int(mc_ptr)
        self.handle = handle

    def ptrs(self, elem_offset: int = 0) -> C.Array:
        b = elem_offset * self.tensor.element_size()
        return ptr_array([p + b for p in self.peer_ptrs])

    def mc(self, elem_offset: int = 0) -> vp:
        return vp(self.mc_ptr + elem_offset * self.tensor.element_size() if self.mc_ptr else 0)


class SymmAllocator:
    """Allocates symmetric buffers for a :class:`parallel.world.World`."""

    def __init__(self, world, use_multicast: bool = True):
        self.world = world
        self.use_multicast = use_multicast
        self._group_name = None
        if world.is_distributed:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            self._symm = symm_mem
            grp = world.group if world.group is not None else dist.group.WORLD
            self._group_name = grp.group_name
            try:
                symm_mem.enable_symm_mem_for_group(self._group_name)
            except Exception:  # noqa: BLE001 - newer torch enables it implicitly
                pass
        self.pads = self.alloc(64, torch.int32, zero=True)   # signal pad for our own barrier protocol

    def alloc(self, numel: int, dtype, zero: bool = True) -> SymmBuffer:
        w = self.world
        numel = max(int(numel), 4)
        if not w.is_distributed:
            t = torch.zeros(numel, dtype=dtype, device=w.device) if zero else torch.empty(numel, dtype=dtype, device=w.device)
            return SymmBuffer(t, [t.data_ptr()])
        sm = self._symm
        t = sm.empty(numel, dtype=dtype, device=w.device)
        if zero:
            t.zero_()
        h = sm.rendezvous(t, self._group_name)
        off = t.data_ptr() - int(h.buffer_ptrs[h.rank])
        peers = [int(p) + off for p in h.buffer_ptrs]
        mc = 0
        if self.use_multicast:
            try:
                mc = int(h.multicast_ptr) or 0
                if mc:
                    mc += off
            except Exception:  # noqa: BLE001
                mc = 0
        torch.cuda.synchronize()
        w.barrier()
        return SymmBuffer(t, peers, mc, h)

    def comm_world(self, use_barrier: bool | None = None) -> CommWorld:
        w = self.world
        cw = CommWorld(w.size, w.rank, int(w.is_distributed if use_barrier is None else use_barrier))
        for i, p in enumerate(self.pads.peer_ptrs):
            cw.signal_pads[i] = p
        return cw


class ScalarReducer:
    """Sum of a few device scalars over the ranks of a world through the owner-computes peer-memory kernel
    (csrc/comm.cu:stat_mean_kernel with scale 1) instead of a library collective: the metric aggregation of the
    data-parallel and federated loops. One instance per world (``for_world``)."""
    _cache: dict = {}

    def __init__(self, world, alloc: "SymmAllocator | None" = None):
        self.world = world
        self.alloc = alloc or SymmAllocator(world)
        self.buf = self.alloc.alloc(64, torch.float32)

    @classmethod
    def for_world(cls, world) -> "ScalarReducer":
        r = cls._cache.get(id(world))
        if r is None or r.world is not world:
            r = cls(world)
            cls._cache[id(world)] = r
        return r

    def sum(self, t: torch.Tensor) -> torch.Tensor:
        from ..ops import _lib
        n = t.numel()
        if n > 64:
            raise ValueError("ScalarReducer: at most 64 values")
        n4 = -(-n // 4) * 4
        self.buf.tensor[:n4].zero_()
        self.buf.tensor[:n].copy_(t.reshape(-1).to(torch.float32))
        cw = self.alloc.comm_world()
        _lib.check(_lib.cuda_lib().idcb_stat_reduce(C.byref(cw), self.buf.ptrs(), self.buf.mc(), C.c_longlong(n4),
                                                    C.c_float(1.0), _lib.stream()))
        _lib.count_launch()
        return self.buf.tensor[:n].to(t.dtype).reshape(t.shape).clone()


class EmulatedWorld:
    """N ranks emulated by one process on one device (tests / single-GPU development)."""

    def __init__(self, n_ranks: int, device="cuda"):
        self.size = n_ranks
        self.device = torch.device(device)

    def alloc(self, numel: int, dtype) -> list:
        """-> list of N SymmBuffers (one per emulated rank) sharing the same pointer table"""
        ts = [torch.zeros(max(numel, 4), dtype=dtype, device=self.device) for _ in range(self.size)]
        ptrs = [t.data_ptr() for t in ts]
        return [SymmBuffer(t, ptrs) for t in ts]

    def comm_world(self, rank: int) -> CommWorld:
        return CommWorld(self.size, rank, 0)
