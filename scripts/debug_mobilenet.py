#!/usr/bin/env python
"""Layer-by-layer comparison of the fused MobileNetV2 pipeline against the reference executor (debug aid).
usage: debug_mobilenet.py [arch] [fine_tune_at]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import torch
from test_gpu_models import _pair

arch = sys.argv[1] if len(sys.argv) > 1 else "mobilenet_v2_tiny"
ft = int(sys.argv[2]) if len(sys.argv) > 2 else 0
B = 64
f, r = _pair(arch, fine_tune_at=ft)
torch.manual_seed(3)
x = torch.rand(B, 50, 50, 3, device="cuda").to(torch.bfloat16).float()
y = (torch.rand(B, device="cuda") > 0.5).float()
f.executor.train_batch(x, y, B)
p = f.executor.plans[B]
rex = r.executor


def ref(name):
    with torch.no_grad():
        stats = r.store.stats.clone()
        out = rex.forward(x, training=True, upto=name).permute(0, 2, 3, 1)
        r.store.stats.copy_(stats)
    return out


def rel(a, b, what):
    a, b = a.float(), b.float()
    rms = ((a - b).pow(2).mean().sqrt() / (b.pow(2).mean().sqrt() + 1e-9)).item()
    print(f"{what:34s} rms_rel {rms:.5f}   |fused| {a.pow(2).mean().sqrt().item():.4f} |ref| {b.pow(2).mean().sqrt().item():.4f}")


rel(p.s_raw, ref("Conv1"), "Conv1")
for b in p.blocks:
    if b.has_expand:
        rel(b.e_raw, ref(b.prefix + "expand"), b.prefix + "expand")
    rel(b.d_raw, ref(b.prefix + "depthwise"), b.prefix + "depthwise")
    rel(b.p_raw, ref(b.prefix + "project"), b.prefix + "project")
    rel(b.out, ref(b.prefix + ("add" if b.add else "project_BN")), b.prefix + ("add" if b.add else "project_BN"))
rel(p.c_raw, ref("Conv_1"), "Conv_1")
