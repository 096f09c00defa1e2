#!/usr/bin/env python
"""Micro-benchmark of the conv kernels at DenseNet/VGG shapes (CUDA events around a graph replay)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C

torch.manual_seed(0)
dev = "cuda"


def bench(fn, iters=20, warm=2):
    """GPU-only time per launch: `iters` launches captured in a CUDA graph, replayed between two events
    (eager ctypes launches cost ~10 us of host time each and would hide the small kernels)."""
    st = torch.cuda.Stream()
    gph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(gph, stream=st):
            for _ in range(iters):
                fn()
    torch.cuda.synchronize()
    gph.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    gph.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) * 1000 / iters


CASES = [  # name, B, H, Cin, Cout, k, pitch
    ("b4 1x1 992->128 M=128", 128, 1, 992, 128, 1),
    ("b4 3x3 128->32 M=128", 128, 1, 128, 32, 3),
    ("b3 1x1 512->128 M=1152", 128, 3, 512, 128, 1),
    ("b3 3x3 128->32 M=1152", 128, 3, 128, 32, 3),
    ("b2 1x1 256->128 M=4608", 128, 6, 256, 128, 1),
    ("b2 3x3 128->32 M=4608", 128, 6, 128, 32, 3),
    ("b1 1x1 128->128 M=21632", 128, 13, 128, 128, 1),
    ("b1 3x3 128->32 M=21632", 128, 13, 128, 32, 3),
    ("vgg 3x3 64->64 M=320000", 128, 50, 64, 64, 3),
    ("vgg 3x3 256->256 M=18432", 128, 12, 256, 256, 3),
    ("vgg 3x3 512->512 M=4608", 128, 6, 512, 512, 3),
]
which = sys.argv[1:] or ["fprop", "dgrad", "wgrad"]
print(f"{'case':28s} " + " ".join(f"{w:>22s}" for w in which))
for name, b, h, cin, cout, k in CASES:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.zeros(b, h, h, cout, device=dev, dtype=torch.bfloat16)
    dx = torch.zeros_like(x)
    dw = torch.zeros(cout, k, k, cin, device=dev)
    sc = torch.ones(cin, device=dev)
    sh = torch.zeros(cin, device=dev)
    s1 = torch.zeros(cout, device=dev)
    s2 = torch.zeros(cout, device=dev)
    flops = 2.0 * b * h * h * cin * cout * k * k
    outs = []
    for wh in which:
        if wh == "fprop":
            t = bench(lambda: C.conv_fprop_raw(g, x, w, y, pro_scale=sc, pro_shift=sh, pro_act=1, stat_sum=s1, stat_sq=s2))
        elif wh == "fprop_plain":
            t = bench(lambda: C.conv_fprop_raw(g, x, w, y))
        elif wh == "dgrad":
            t = bench(lambda: C.conv_dgrad_raw(g, y, w, dx))
        elif wh == "dgrad2":
            G = torch.zeros(b, h, h, cin, device=dev)
            t = bench(lambda: C.conv_dgrad_raw(g, y, w, None, dx_pitch=cin, mode=2, aux=x, bwd_scale=sc, bwd_shift=sh, bwd_act=1, gacc=G, stat_a=sc.clone(), stat_b=sh.clone()))
        else:
            t = bench(lambda: C.conv_wgrad_raw(g, x, y, dw, pro_scale=sc, pro_shift=sh, pro_act=1))
        outs.append(f"{t:8.1f}us {flops / t / 1e6:8.1f}TF/s")
    print(f"{name:28s} " + " ".join(f"{o:>22s}" for o in outs))
