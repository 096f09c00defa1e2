"""Process-group plumbing: one process per GPU (or per CPU replica), ``torch.distributed`` for
bootstrap, NCCL (GPU) or gloo (CPU) as the library transport.

The reference is single-process / in-graph replicated (``tf.distribute.MirroredStrategy`` drives N
GPUs from one Python thread: /root/reference/dist_model_tf_mobile.py:115-117). With sub-millisecond
steps on B200 a single host thread would be the bottleneck, so here every replica is its own process
launched by ``torchrun`` / ``python -m torch.distributed.run``; ``World`` reads RANK / LOCAL_RANK /
WORLD_SIZE / MASTER_ADDR / MASTER_PORT from the environment.
"""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist


class World:
    _default: "World | None" = None

    def __init__(self, rank: int = 0, size: int = 1, local_rank: int = 0, device: str | None = None,
                 group=None):
        self.rank, self.size, self.local_rank = rank, size, local_rank
        if device is None:
            device = f"cuda:{local_rank}" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device)
        self.group = group

    @property
    def is_distributed(self) -> bool:
        return self.size > 1

    @property
    def backend(self) -> str:
        return dist.get_backend(self.group) if self.is_distributed else "none"

    @classmethod
    def from_env(cls, device: str | None = None, timeout_s: int = 600) -> "World":
        """Join the job described by the torchrun environment (or a single-process world)."""
        if cls._default is not None:
            return cls._default
        size = int(os.environ.get("WORLD_SIZE", "1"))
        rank = int(os.environ.get("RANK", "0"))
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        use_cuda = torch.cuda.is_available() and (device is None or str(device).startswith("cuda"))
        if use_cuda:
            torch.cuda.set_device(local_rank)
        if size > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            kw = {}
            if use_cuda:
                kw["device_id"] = torch.device(f"cuda:{local_rank}")
            dist.init_process_group("nccl" if use_cuda else "gloo", rank=rank, world_size=size,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)
        w = cls(rank, size, local_rank, (f"cuda:{local_rank}" if use_cuda else "cpu"),
                dist.group.WORLD if size > 1 else None)
        cls._default = w
        return w

    @classmethod
    def single(cls, device: str | None = None) -> "World":
        return cls(0, 1, 0, device)

    # ---- collectives over the library transport (NCCL / gloo) ------------------------------
    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.is_distributed:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def broadcast(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.is_distributed:
            dist.broadcast(t, src=src, group=self.group)
        return t

    def reduce_to(self, t: torch.Tensor, dst: int = 0) -> torch.Tensor:
        if self.is_distributed:
            dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_gather_object(self, obj):
        if not self.is_distributed:
            return [obj]
        out = [None] * self.size
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self) -> None:
        if self.is_distributed:
            if self.device.type == "cuda":
                dist.barrier(group=self.group, device_ids=[self.device.index])
            else:
                dist.barrier(group=self.group)

    def shutdown(self) -> None:
        if dist.is_initialized():
            # a rank that tears the transport down while a peer is still inside its last collective makes
            # that peer abort (gloo: SIGABRT); leave together
            try:
                self.barrier()
            except Exception:  # noqa: BLE001 - a peer that already died must not turn shutdown into a hang report
                pass
            dist.destroy_process_group()
        World._default = None
