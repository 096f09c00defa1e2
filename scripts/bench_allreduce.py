#!/usr/bin/env python
"""Fused all-reduce + optimizer kernel vs NCCL all-reduce + optimizer launch, on the gradient sizes of the
three models. Launch with torchrun on N >= 2 GPUs. Device-timed, max over ranks; reports the
achieved fraction of the NVLink roofline S / link_bw (link_bw = 770 GB/s measured peer copy, 900 nominal)."""
import json
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist
from idc_models_b200.optim.optimizers import RMSprop
from idc_models_b200.parallel.fused_comm import FusedComm
from idc_models_b200.parallel.world import World


class Store:
    def __init__(self, n, dev):
        self.master = torch.randn(n, device=dev)
        self.grad = torch.randn(n, device=dev)
        self.shadow = self.master.to(torch.bfloat16)
        self.version = self.ext_version = 0

    def touch(self, external=False):
        self.version += 1


world = World.from_env()
dev = world.device
res = []
for name, n in (("densenet121 (fp32 grads)", 6954881), ("vgg16 (fp32 grads)", 14715201), ("mobilenet_v2", 2225153)):
    n = -(-n // 256) * 256
    for use_mc in (True, False):
        st = Store(n, dev)
        fc = FusedComm(world, st, use_multicast=use_mc)
        if use_mc and not fc.uses_multicast:
            continue
        opt = RMSprop(1e-4).bind(st, [(0, n)])
        for _ in range(5):
            fc.allreduce_apply(opt, 1.0)
        torch.cuda.synchronize(); world.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
        for s, e in ev:
            s.record(); fc.allreduce_apply(opt, 1.0); e.record()
        torch.cuda.synchronize()
        t = torch.tensor([sorted(s.elapsed_time(e) for s, e in ev)[len(ev) // 2]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res.append((name, "fused+mc" if use_mc else "fused p2p", n, float(t.item()) * 1e3))
    # NCCL baseline: all_reduce + optimizer kernel
    st = Store(n, dev)
    opt = RMSprop(1e-4).bind(st, [(0, n)])
    for _ in range(5):
        dist.all_reduce(st.grad); opt.apply(1.0)
    torch.cuda.synchronize(); world.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for s, e in ev:
        s.record(); dist.all_reduce(st.grad); opt.apply(1.0); e.record()
    torch.cuda.synchronize()
    t = torch.tensor([sorted(s.elapsed_time(e) for s, e in ev)[len(ev) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res.append((name, "nccl allreduce + optimizer", n, float(t.item()) * 1e3))
if world.rank == 0:
    N = world.size
    for name, kind, n, us in res:
        # bytes that must cross this GPU's NVLink per direction: (N-1)/N of the fp32 gradient in, and
        # (N-1)/N of fp32 weights + bf16 shadow out
        s_in = 4 * n * (N - 1) / N
        s_out = 6 * n * (N - 1) / N
        roof_us = max(s_in, s_out) / 770e9 * 1e6
        print(json.dumps({"case": name, "path": kind, "n_gpus": N, "elements": n, "us": round(us, 1),
                          "roofline_us_770GBs": round(roof_us, 1), "fraction_of_roofline": round(roof_us / us, 3)}))
world.barrier()
world.shutdown()
