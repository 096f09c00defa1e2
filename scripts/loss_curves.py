#!/usr/bin/env python
"""Loss curves of the fused pipeline and of the bf16-emulating reference executor on the same data (DenseNet-121,
50 steps), for RMSprop and SGD: how far two numerically equivalent implementations drift apart step by step."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import RMSprop, SGD

for optname in ("rmsprop", "sgd"):
    torch.manual_seed(1)
    nets = []
    for ex in ("fused", "reference_bf16", "reference"):
        net = build_transfer_model("densenet121", (50, 50, 3), 1, device="cuda", seed=0, executor=ex)
        net.base.trainable = True
        opt = RMSprop(lr=1e-4) if optname == "rmsprop" else SGD(lr=1e-2)
        net.compile(optimizer=opt, loss="bce", metrics=["accuracy"])
        nets.append(net)
    for n in nets[1:]:
        n.store.master.copy_(nets[0].store.master); n.store.stats.copy_(nets[0].store.stats)
    g = torch.Generator(device="cpu").manual_seed(9)
    data = []
    for _ in range(5):
        x = torch.rand(32, 50, 50, 3, generator=g).cuda(); y = (torch.rand(32, generator=g) > 0.5).float().cuda()
        x[y > 0.5] *= 0.8
        data.append((x, y))
    for net, name in zip(nets, ("fused", "ref_bf16", "ref_fp32")):
        losses = []
        for step in range(50):
            x, y = data[step % 5]
            losses.append(float(net.train_on_batch(x, y)["stats"][0].item()) / 32)
        print(optname, name, " ".join(f"{v:.4f}" for v in losses))
