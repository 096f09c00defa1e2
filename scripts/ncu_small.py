#!/usr/bin/env python
"""The DenseNet-121 dense-layer kernels at the block-3 shapes (M = 128 x 3 x 3 = 1152 rows, latency-bound), each
launched twice, as a target for `ncu --set full --import-source on`: 1x1 fprop with the BatchNorm prologue, 3x3
fprop with prologue, 3x3 dgrad (mode 2 + BatchNorm tail), 1x1 dgrad (mode 2 into G), both weight gradients."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C

dev = "cuda"
b, h = 128, int(sys.argv[1]) if len(sys.argv) > 1 else 3
cin = 512
g1 = C.geom(h, h, cin, 128, 1)
g3 = C.geom(h, h, 128, 32, 3)
buf = torch.randn(b, h, h, 1024, device=dev).to(torch.bfloat16)
y1 = torch.randn(b, h, h, 128, device=dev).to(torch.bfloat16)
w1 = (torch.randn(128, 1, 1, cin, device=dev) * 0.05).to(torch.bfloat16)
w3 = (torch.randn(32, 3, 3, 128, device=dev) * 0.05).to(torch.bfloat16)
dy3 = torch.randn(b, h, h, 32, device=dev).to(torch.bfloat16)
dy1 = torch.zeros(b, h, h, 128, device=dev, dtype=torch.bfloat16)
dpre = torch.zeros_like(dy1)
G = torch.zeros(b, h, h, 1024, device=dev)
sc1, sh1 = torch.ones(cin, device=dev), torch.zeros(cin, device=dev)
sc3, sh3 = torch.ones(128, device=dev), torch.zeros(128, device=dev)
mean, rstd = torch.zeros(128, device=dev), torch.ones(128, device=dev)
s = [torch.zeros(1024, device=dev) for _ in range(8)]
dw1 = torch.zeros(128, 1, 1, cin, device=dev)
dw3 = torch.zeros(32, 3, 3, 128, device=dev)
ctr = torch.zeros(64, device=dev)
for rep in range(2):
    C.conv_fprop_raw(g1, buf, w1, y1, pro_scale=sc1, pro_shift=sh1, pro_act=1, stat_sum=s[0][:128], stat_sq=s[1][:128])
    C.conv_fprop_raw(g3, y1, w3, buf, y_coff=cin, pro_scale=sc3, pro_shift=sh3, pro_act=1, stat_sum=s[2][:32],
                     stat_sq=s[3][:32])
    ctr.zero_()
    C.conv_dgrad_raw(g3, dy3, w3, dpre, mode=2, aux=y1, bwd_scale=sc3, bwd_shift=sh3, bwd_act=1,
                     stat_a=s[4][:128], stat_b=s[5][:128],
                     tail=C.bn_tail(dy1, mean, rstd, b * h * h, counter=ctr))
    C.conv_dgrad_raw(g1, dy1, w1, None, dx_pitch=1024, mode=2, aux=buf, bwd_scale=sc1, bwd_shift=sh1, bwd_act=1,
                     gacc=G, stat_a=s[6][:cin], stat_b=s[7][:cin])
    C.conv_wgrad_raw(g3, y1, dy3, dw3, pro_scale=sc3, pro_shift=sh3, pro_act=1)
    C.conv_wgrad_raw(g1, buf, dy1, dw1, pro_scale=sc1, pro_shift=sh1, pro_act=1)
torch.cuda.synchronize()
print("ok")
