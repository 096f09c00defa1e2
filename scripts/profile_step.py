#!/usr/bin/env python
"""A few eager (no CUDA graph) training steps of the flagship model, for `ncu` launch lists:

    python scripts/profile_step.py --steps 1 && \
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
        python scripts/profile_step.py --steps 1
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from idc_models_b200.models.applications import build_transfer_model  # noqa: E402
from idc_models_b200.optim.optimizers import RMSprop  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="densenet121")
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--warmup", type=int, default=1)
ap.add_argument("--fine-tune-at", type=int, default=0)
a = ap.parse_args()
m = build_transfer_model(a.model, (50, 50, 3), 1, device="cuda")
m.base.trainable = True
for l in m.base.layers[:a.fine_tune_at]:
    l.trainable = False
m.compile(optimizer=RMSprop(lr=1e-4), loss="bce", metrics=["accuracy"])
x = torch.rand(a.batch, 50, 50, 3, device="cuda")
y = (torch.rand(a.batch, device="cuda") > 0.5).float()
for i in range(a.warmup + a.steps):
    if i == a.warmup:
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_push("timed")
    m.train_on_batch(x, y)
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print("ok")
