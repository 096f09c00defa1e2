"""Multi-process NVLink tests (>= 2 GPUs on the box; skipped otherwise): the fused all-reduce+optimizer
step against NCCL all-reduce + separate optimizer, replicas bit-identical, and the FedAvg kernel across
ranks. Run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0

WORKER = r'''
import sys, json, torch
sys.path.insert(0, %r)
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import SGD
from idc_models_b200.parallel.strategy import MirroredStrategy
from idc_models_b200.parallel.world import World
world = World.from_env()
res = {}
for comm in ("nccl", "fused"):
    strategy = MirroredStrategy(world=world, comm=comm)
    with strategy.scope():
        m = build_transfer_model("densenet_tiny", device=world.device, seed=0)
        m.base.trainable = True
        m.compile(optimizer=SGD(learning_rate=1e-2), loss="bce", metrics=["accuracy"])   # linear in g: tight check
    g = torch.Generator().manual_seed(0)
    gb = 32 * world.size
    x = torch.rand(gb, 50, 50, 3, generator=g).to(world.device); y = (torch.rand(gb, generator=g) > 0.5).float().to(world.device)
    for _ in range(3):
        m.train_on_batch(x, y)
    torch.cuda.synchronize()
    mc = bool(strategy._fused.uses_multicast) if comm == "fused" else None
    res[comm] = (m.store.master.double().sum().item(), m.store.master.double().abs().sum().item(), mc)
outs = world.all_gather_object(res)
if world.rank == 0:
    print("RESULT " + json.dumps(outs))
world.barrier()
world.shutdown()
''' % str(ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_fused_allreduce_matches_nccl_path_across_processes():
    script = ROOT / "build" / "_multi_worker.py"
    script.parent.mkdir(exist_ok=True)
    script.write_text(WORKER)
    n = min(NGPU, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    outs = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    print(outs)
    for o in outs:
        assert o["fused"][:2] == outs[0]["fused"][:2], "fused replicas are not bit-identical"
        # NCCL ring order vs owner-computes order differ in fp32 rounding only
        assert abs(o["fused"][0] - o["nccl"][0]) < 1e-4 * max(1.0, abs(o["nccl"][1]))
    print("multicast (NVLS) used by the fused path:", outs[0]["fused"][2])
