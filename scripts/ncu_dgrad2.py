#!/usr/bin/env python
"""Launches the BN-backward (mode 2) data-gradient kernel at the DenseNet block-1 shapes (target for `ncu --set full`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C

dev = "cuda"
for (b, h, cin, cout, k) in [(128, 13, 128, 128, 1), (128, 13, 128, 32, 3)]:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.randn(b, h, h, cout, device=dev).to(torch.bfloat16)
    sc = torch.ones(cin, device=dev); sh = torch.zeros(cin, device=dev)
    G = torch.zeros(b, h, h, cin, device=dev)
    sa, sb = sc.clone(), sh.clone()
    for _ in range(3):
        C.conv_dgrad_raw(g, y, w, None, dx_pitch=cin, mode=2, aux=x, bwd_scale=sc, bwd_shift=sh, bwd_act=1, gacc=G,
                         stat_a=sa, stat_b=sb)
    torch.cuda.synchronize()
print("ok")
