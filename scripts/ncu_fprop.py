#!/usr/bin/env python
"""Launches the forward conv kernels of the small DenseNet shapes once each (target for `ncu --set full`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C

dev = "cuda"
for (b, h, cin, cout, k) in [(128, 3, 512, 128, 1), (128, 3, 128, 32, 3), (128, 1, 992, 128, 1), (128, 13, 128, 32, 3)]:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.zeros(b, h, h, cout, device=dev, dtype=torch.bfloat16)
    sc = torch.ones(cin, device=dev); sh = torch.zeros(cin, device=dev)
    s1 = torch.zeros(cout, device=dev); s2 = torch.zeros(cout, device=dev)
    for _ in range(2):
        C.conv_fprop_raw(g, x, w, y)
        C.conv_fprop_raw(g, x, w, y, pro_scale=sc, pro_shift=sh, pro_act=1, stat_sum=s1, stat_sq=s2)
    torch.cuda.synchronize()
print("ok")
