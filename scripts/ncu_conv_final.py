#!/usr/bin/env python
"""One launch each of the three conv directions at the DenseNet block-1 3x3 shape (M = 21632, 128 -> 32) and the
block-3 bottleneck 1x1 (M = 1152, 512 -> 128), with the fused BatchNorm prologue / backward epilogue the model uses.
Target for `ncu --set full`; the summary goes to profiles/conv_kernels_ncu.md."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C

dev = "cuda"
for (b, h, cin, cout, k) in [(128, 13, 128, 32, 3), (128, 3, 512, 128, 1)]:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.randn(b, h, h, cout, device=dev).to(torch.bfloat16)
    dw = torch.zeros(cout, k, k, cin, device=dev)
    sc = torch.ones(cin, device=dev); sh = torch.zeros(cin, device=dev)
    s1 = torch.zeros(cout, device=dev); s2 = torch.zeros(cout, device=dev)
    G = torch.zeros(b, h, h, cin, device=dev)
    sa, sb = sc.clone(), sh.clone()
    for _ in range(2):
        C.conv_fprop_raw(g, x, w, y, pro_scale=sc, pro_shift=sh, pro_act=1, stat_sum=s1, stat_sq=s2)
        C.conv_dgrad_raw(g, y, w, None, dx_pitch=cin, mode=2, aux=x, bwd_scale=sc, bwd_shift=sh, bwd_act=1, gacc=G,
                         stat_a=sa, stat_b=sb)
        C.conv_wgrad_raw(g, x, y, dw, pro_scale=sc, pro_shift=sh, pro_act=1)
    torch.cuda.synchronize()
print("ok")
