#!/usr/bin/env python
"""Per-phase clock stamps of one conv CTA (profiling aid) + graph-replayed GPU-only timings."""
import ctypes, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C, _lib

dev = "cuda"
lib = _lib.cuda_lib()
dbg = torch.zeros(16, dtype=torch.int64, device=dev)
names = ["entry", "setup", "A0", "prod_done", "mma_first", "mma_issued", "acc_ready", "pre_sync", "post_sync", "dealloc", "t0_arrived", "t224_arrived", "mma_woke", "", "", ""]


def graph_time(fn, iters=20):
    gph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(gph, stream=s):
            for _ in range(iters):
                fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gph.replay(); torch.cuda.synchronize()
    e0.record(); gph.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / iters


for (b, h, cin, cout, k) in [(128, 1, 992, 128, 1), (128, 1, 128, 32, 3), (128, 3, 512, 128, 1), (128, 3, 128, 32, 3),
                             (128, 13, 128, 128, 1), (128, 13, 128, 32, 3)]:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.zeros(b, h, h, cout, device=dev, dtype=torch.bfloat16)
    dx = torch.zeros_like(x)
    dw = torch.zeros(cout, k, k, cin, device=dev)
    sc = torch.ones(cin, device=dev); sh = torch.zeros(cin, device=dev)
    s1 = torch.zeros(cout, device=dev); s2 = torch.zeros(cout, device=dev)
    G = torch.zeros(b, h, h, cin, device=dev)
    sa, sb = sc.clone(), sh.clone()
    fns = {
        "fprop": lambda: C.conv_fprop_raw(g, x, w, y),
        "fprop+bn": lambda: C.conv_fprop_raw(g, x, w, y, pro_scale=sc, pro_shift=sh, pro_act=1, stat_sum=s1, stat_sq=s2),
        "dgrad": lambda: C.conv_dgrad_raw(g, y, w, dx),
        "dgrad2": lambda: C.conv_dgrad_raw(g, y, w, None, dx_pitch=cin, mode=2, aux=x, bwd_scale=sc, bwd_shift=sh,
                                           bwd_act=1, gacc=G, stat_a=sa, stat_b=sb),
        "wgrad+bn": lambda: C.conv_wgrad_raw(g, x, y, dw, pro_scale=sc, pro_shift=sh, pro_act=1),
    }
    for label, fn in fns.items():
        fn(); torch.cuda.synchronize()
        lib.idcb_conv_set_debug(ctypes.c_void_p(dbg.data_ptr()))
        dbg.zero_()
        fn()
        torch.cuda.synchronize()
        lib.idcb_conv_set_debug(ctypes.c_void_p(0))
        t = dbg.tolist()
        us = graph_time(fn)
        print(f"{label:9s} {b}x{h}x{h} {cin}->{cout} k{k}: {us:6.2f} us/launch | "
              + ", ".join(f"{names[i]}={t[i]-t[0]}" for i in range(1, 13) if t[i]))
# trivial kernel for reference
from idc_models_b200.ops import kernels as K
a = torch.zeros(64, device=dev); bb = torch.zeros(64, device=dev)
print(f"bn_fold (trivial kernel) graph replay: {graph_time(lambda: K.bn_fold(a, a, a, a + 1, 1e-3, bb, bb)):.2f} us per launch")
