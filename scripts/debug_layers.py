#!/usr/bin/env python
"""Layer-by-layer comparison of the fused DenseNet pipeline against the reference executor in training
mode (debug aid): prints the relative error of each stored activation."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
from test_gpu_models import _pair

arch = sys.argv[1] if len(sys.argv) > 1 else "densenet_tiny"
B = 64
f, r = _pair(arch)
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
    e = (a - b).abs().max().item()
    s = b.abs().max().item() + 1e-9
    rms = ((a - b).pow(2).mean().sqrt() / (b.pow(2).mean().sqrt() + 1e-9)).item()
    print(f"{what:34s} max_err/max {e / s:.4f}  rms_rel {rms:.5f}")


rel(p.s0, ref("conv1/conv"), "conv1/conv")
b0 = p.blocks[0]
rel(b0.buf[..., :64], ref("pool1"), "pool1")
for bi, blk in enumerate(p.blocks):
    for dl in blk.layers:
        rel(dl.y1, ref(dl.name + "_1_conv"), dl.name + "_1_conv")
        rel(blk.buf[..., dl.cin:dl.cin + 32], ref(dl.name + "_2_conv"), dl.name + "_2_conv")
    if hasattr(blk, "t"):
        rel(blk.t, ref(f"pool{bi + 2}_conv"), f"pool{bi + 2}_conv")
        nxt = p.blocks[bi + 1]
        rel(nxt.buf[..., :nxt.c0], ref(f"pool{bi + 2}_pool"), f"pool{bi + 2}_pool")
lb = p.blocks[-1]
feat = torch.relu(lb.buf.float() * lb.bn_t.scale + lb.bn_t.shift)
rel(feat, ref("relu"), "final relu (from fused scale/shift)")
