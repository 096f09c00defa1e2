#!/usr/bin/env python
"""Fused all-reduce + optimizer kernel vs NCCL all-reduce + optimizer launch, on the gradient sizes of the
three models. Launch with torchrun on N >= 2 GPUs. Device-timed (graph replay), max over ranks; reports the
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
REPS = 10


def graph_time(fn) -> float:
    """device microseconds per call: REPS calls captured in a CUDA graph and replayed between two events (host launch
    cost, ~25 us per eager call through Python, would otherwise sit inside the timed region); max over ranks"""
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); world.barrier()
    st_ = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st_):
        with torch.cuda.graph(g, stream=st_):
            for _ in range(REPS):
                fn()
    torch.cuda.synchronize(); world.barrier()
    g.replay(); torch.cuda.synchronize(); world.barrier()
    times = []
    for _ in range(5):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        world.barrier(); torch.cuda.synchronize()
        s.record(); g.replay(); e.record()
        torch.cuda.synchronize()
        times.append(s.elapsed_time(e) * 1e3 / REPS)
    t = torch.tensor([sorted(times)[len(times) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
import os
CASES = (("densenet121 (fp32 grads)", 6954881), ("vgg16 (fp32 grads)", 14715201), ("mobilenet_v2", 2225153), ("tiny: fixed cost (launch + 2 barriers)", 4096))
if os.environ.get("IDCB_AR_CASES"):
    CASES = tuple(c for c in CASES if any(k in c[0] for k in os.environ["IDCB_AR_CASES"].split(",")))
for name, n in CASES:
    n = -(-n // 256) * 256
    for use_mc in (True, False):
        st = Store(n, dev)
        fc = FusedComm(world, st, use_multicast=use_mc)
        if use_mc and not fc.uses_multicast:
            continue
        opt = RMSprop(1e-4).bind(st, [(0, n)])
        us = graph_time(lambda: fc.allreduce_apply(opt, 1.0))
        res.append((name, "fused+mc" if use_mc else "fused p2p", n, us))
    # NCCL baseline: all_reduce + optimizer kernel
    st = Store(n, dev)
    opt = RMSprop(1e-4).bind(st, [(0, n)])
    def nccl_step():
        dist.all_reduce(st.grad)
        opt.apply(1.0)
    us = graph_time(nccl_step)
    res.append((name, "nccl allreduce + optimizer", n, us))
if world.rank == 0:
    N = world.size
    for name, kind, n, us in res:
        # bytes that must cross this GPU's NVLink per direction: (N-1)/N of the fp32 gradient in, and
        # (N-1)/N of fp32 weights + bf16 shadow out
        s_in = 4 * n * (N - 1) / N
        s_out = 6 * n * (N - 1) / N
        roof_us = max(s_in, s_out) / 770e9 * 1e6          # measured peer-copy bandwidth per direction (B200_PROFILING.md)
        roof900 = max(s_in, s_out) / 900e9 * 1e6          # nominal NVLink 5 figure
        print(json.dumps({"case": name, "path": kind, "n_gpus": N, "elements": n, "us": round(us, 1),
                          "roofline_us_770GBs": round(roof_us, 1), "fraction_of_roofline": round(roof_us / us, 3),
                          "roofline_us_900GBs": round(roof900, 1), "fraction_of_900": round(roof900 / us, 3)}))
world.barrier()
world.shutdown()
