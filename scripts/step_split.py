#!/usr/bin/env python
"""Where a fused training step spends its time: forward alone (evaluation graph), full step with and without the
weight-gradient side stream, all as CUDA-graph replays timed with events."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import RMSprop

model = sys.argv[1] if len(sys.argv) > 1 else "densenet121"


def timed(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


x = torch.rand(128, 50, 50, 3, device="cuda")
y = (torch.rand(128, device="cuda") > 0.5).float()
for side in (True, False):
    m = build_transfer_model(model, (50, 50, 3), 1, device="cuda")
    m.base.trainable = True
    m.compile(optimizer=RMSprop(lr=1e-4), loss="bce", metrics=["accuracy"])
    ex = m.executor
    ex.use_graphs = True
    ex.wgrad_side_stream = side
    t_step = timed(lambda: (ex.train_batch(x, y, 128), m.optimizer.apply(1.0)))
    t_eval = timed(lambda: ex.eval_batch(x, y))
    print(f"{model} side_stream={side}: train step {t_step:.3f} ms, eval forward {t_eval:.3f} ms, step - forward {t_step - t_eval:.3f} ms")
