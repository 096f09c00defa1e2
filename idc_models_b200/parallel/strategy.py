"""Distribution strategies with the ``tf.distribute`` surface the reference scripts use
(``strategy.scope()``, ``strategy.num_replicas_in_sync``), re-designed as one process per replica.

* :class:`MirroredStrategy` — synchronous data parallelism (SURVEY.md 2.B B12, 2.D;
  /root/reference/dist_model_tf_dense.py:22, dist_model_tf_mobile.py:115, dist_model_tf_vgg.py:115):
  identical variables on every replica, the global batch split across replicas, ONE packed gradient
  sum per step followed by an identical update everywhere. ``comm='fused'`` runs the hand-written
  NVLink kernel that fuses the reduction with the optimizer update (parallel/fused_comm.py);
  ``comm='nccl'``/``'gloo'`` call the library all-reduce and then the optimizer (the baseline).
* :class:`CentralStorageStrategy` — variables and optimizer state owned by ONE replica
  (dist_model_tf_dense.py:24 behind ``use_mirror=False``): gradients are reduced to
  the owner, the owner updates, new values are broadcast.
* default (no scope) — a single replica.
"""
from __future__ import annotations

import contextlib
import threading

import torch

from .world import World

_tls = threading.local()


def current_strategy() -> "Strategy":
    s = getattr(_tls, "strategy", None)
    return s if s is not None else _DEFAULT


class Strategy:
    name = "default"

    def __init__(self, world: World | None = None, comm: str = "auto"):
        self.world = world or World.single()
        if comm == "auto":
            comm = "nccl" if self.world.device.type == "cuda" else "gloo"
        self.comm = comm
        self._fused = None

    @property
    def num_replicas_in_sync(self) -> int:
        return self.world.size

    @property
    def replica_id(self) -> int:
        return self.world.rank

    @contextlib.contextmanager
    def scope(self):
        prev = getattr(_tls, "strategy", None)
        _tls.strategy = self
        try:
            yield self
        finally:
            _tls.strategy = prev

    # ---- batch splitting (global batch -> this replica's slice) ------------------------------
    def local_slice(self, n: int) -> slice:
        per = -(-n // self.world.size)
        lo = min(self.world.rank * per, n)
        return slice(lo, min(lo + per, n))

    # ---- variable creation (C3): make replicas identical -------------------------------------
    def sync_variables(self, store) -> None:
        if self.world.is_distributed:
            self.world.broadcast(store.master, 0)
            self.world.broadcast(store.stats, 0)
            store.refresh_shadow()

    # ---- per-step gradient reduction + update (C1 + K13) --------------------------------------
    def reduce_and_apply(self, store, optimizer, grad_scale: float = 1.0, scalars=None):
        """Sum gradients over replicas and apply the optimizer on every replica. ``scalars`` (the step's
        device-side loss-sum / correct / count slots) are summed over replicas on the way where the transport
        can carry them (the fused kernel's barrier); the global sums are returned, else None."""
        w = self.world
        if not w.is_distributed:
            optimizer.apply(grad_scale)
            return None
        if self.comm == "fused":
            sched = getattr(self, "_overlap", None)
            if sched is not None and sched.optimizer is optimizer:
                # the kernels were launched from inside the backward pass (or are launched now, by a rank whose
                # local batch was empty); only the host-side step bookkeeping is left
                if not sched.consumed:
                    sched.run_all()
                sched.consumed = False
                optimizer.iterations += 1
                store.touch()
                return sched.fused.scal_out[:3] if scalars is not None else None
            out = self._fused_comm(store).allreduce_apply(optimizer, grad_scale, scalars=scalars)
            return out[:3] if out is not None else None
        for b, e in optimizer.ranges:
            w.all_reduce_sum(store.grad[b:e])
        optimizer.apply(grad_scale)
        return None

    def _fused_comm(self, store):
        if self._fused is None or self._fused.store is not store:
            from .fused_comm import FusedComm
            self._fused = FusedComm(self.world, store)
        return self._fused

    overlap = True      # fused path only: launch gradient buckets from inside the backward pass when possible
    overlap_min_bucket = 1 << 19    # elements; smaller announcements are merged (launch + barrier latency per bucket)

    def after_compile(self, net) -> None:
        self._overlap = None
        if self.overlap and self.comm == "fused":
            self.enable_overlap(net, self.overlap_min_bucket)

    def enable_overlap(self, net, min_bucket: int = 1 << 19) -> bool:
        """Overlap the gradient all-reduce with the backward pass of ``net`` (compiled, fused executor).
        Returns False (and changes nothing) where that is not possible."""
        from .fused_comm import OverlapSchedule
        ex = net.executor
        if self.comm != "fused" or not self.world.is_distributed or not hasattr(ex, "grad_ready_offsets"):
            return False
        try:
            sched = OverlapSchedule(self._fused_comm(net.store), net.optimizer, ex, 1.0, min_bucket)
        except ValueError:
            return False
        self._overlap = sched
        ex.grad_hook = sched
        return True

    def prepare_store(self, store) -> None:
        """Called by ``compile`` before the optimizer and executor bind to the flat buffers: the fused
        path re-seats them in symmetric (peer-mapped) memory."""
        if self.comm == "fused" and self.world.is_distributed:
            self._fused_comm(store)

    # ---- scalar metric reduction (C2) -----------------------------------------------------------
    def reduce_scalars(self, t: torch.Tensor) -> torch.Tensor:
        if (self.comm == "fused" and self._fused is not None and self.world.is_distributed and t.is_cuda
                and t.numel() <= 64 and t.dtype in (torch.float32, torch.float16, torch.bfloat16)):
            return self._fused.reduce_scalars(t)
        return self.world.all_reduce_sum(t)

    # ---- BN moving statistics: MEAN over replicas when read (C4) -------------------------------
    def aggregate_moving_stats(self, store) -> None:
        if self.world.is_distributed and store.n_stats:
            if self.comm == "fused" and self._fused is not None and self._fused.store is store:
                self._fused.stat_mean()
                return
            self.world.all_reduce_sum(store.stats)
            store.stats.div_(self.world.size)


class MirroredStrategy(Strategy):
    name = "mirrored"

    def __init__(self, devices=None, world: World | None = None, comm: str = "auto"):
        """``devices`` is accepted for signature parity (dist_model_tf_dense.py:22); the number of
        replicas is the number of launched processes (torchrun ``--nproc-per-node``)."""
        super().__init__(world or World.from_env(), comm)
        self.devices = devices


class CentralStorageStrategy(Strategy):
    name = "central_storage"
    overlap = False      # one owner updates everything after the backward pass; nothing is launched from inside it

    def __init__(self, compute_devices=None, parameter_device=None, world: World | None = None,
                 comm: str = "auto", owner: int = 0):
        super().__init__(world or World.from_env(), comm)
        self.owner = owner

    def reduce_and_apply(self, store, optimizer, grad_scale: float = 1.0, scalars=None):
        w = self.world
        if not w.is_distributed:
            optimizer.apply(grad_scale)
            return None
        if self.comm == "fused":
            # same kernel, one owner: reduce-to-owner + update + broadcast over NVLink
            out = self._fused_comm(store).allreduce_apply(optimizer, grad_scale, owner=self.owner, scalars=scalars)
            return out[:3] if out is not None else None
        # gradients -> owner ; owner applies ; fresh variables -> everyone
        for b, e in optimizer.ranges:
            w.reduce_to(store.grad[b:e], self.owner)
        if w.rank == self.owner:
            optimizer.apply(grad_scale)
        else:
            for b, e in optimizer.ranges:
                store.grad[b:e].zero_()
        for b, e in optimizer.ranges:
            w.broadcast(store.master[b:e], self.owner)
        if store.shadow is not None:
            for b, e in optimizer.ranges:
                store.shadow[b:e].copy_(store.master[b:e])


_DEFAULT = Strategy(World.single("cpu"), comm="none")


def get_strategy() -> Strategy:
    return current_strategy()


class FederatedStrategy(Strategy):
    """Clients-to-GPUs placement for the federated workloads: every rank trains its own clients
    independently (no data parallelism inside a client); ranks meet only in the aggregation kernels
    (fed/fedavg.py, fed/secure.py). SURVEY.md 2.D "Federated (client-parallel simulation)"."""
    name = "federated"

    def __init__(self, world: World | None = None):
        super().__init__(world or World.from_env(), comm="none")

    def local_slice(self, n: int) -> slice:
        return slice(0, n)

    def reduce_and_apply(self, store, optimizer, grad_scale: float = 1.0, scalars=None):
        optimizer.apply(grad_scale)
        return None

    def reduce_scalars(self, t):
        return t

    def aggregate_moving_stats(self, store) -> None:
        return None
