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
from idc_models_b200.parallel.strategy import CentralStorageStrategy, MirroredStrategy
from idc_models_b200.parallel.world import World
world = World.from_env()
res = {}
for comm in ("nccl", "fused", "fused_overlap", "fused_overlap_graph", "nccl_central", "fused_central"):
    cls = CentralStorageStrategy if comm.endswith("central") else MirroredStrategy
    strategy = cls(world=world, comm=comm.split("_")[0])
    strategy.overlap = "overlap" in comm
    strategy.overlap_min_bucket = 4096
    with strategy.scope():
        m = build_transfer_model("densenet_tiny", device=world.device, seed=0)
        m.base.trainable = True
        m.compile(optimizer=SGD(learning_rate=1e-2), loss="bce", metrics=["accuracy"])   # linear in g: tight check
    m.executor.use_graphs = comm.endswith("graph")
    if "overlap" in comm:
        assert strategy._overlap is not None and len(strategy._overlap.plan) >= 2, "no overlap schedule was built"
    g = torch.Generator().manual_seed(0)
    gb = 32 * world.size
    x = torch.rand(gb, 50, 50, 3, generator=g).to(world.device); y = (torch.rand(gb, generator=g) > 0.5).float().to(world.device)
    for _ in range(3):
        m.train_on_batch(x, y)
    torch.cuda.synchronize()
    mc = bool(strategy._fused.uses_multicast) if comm.startswith("fused") else None
    if comm == "fused":
        # BN moving statistics differ per replica (own batches); on read they become the mean over replicas
        mine = m.store.stats.detach().cpu().clone()
        allst = world.all_gather_object(mine)
        want = torch.stack(allst).double().mean(0).float()
        strategy.aggregate_moving_stats(m.store)
        torch.cuda.synchronize()
        res["stat_mean_err"] = (m.store.stats.cpu() - want).abs().max().item()
        res["stat_spread"] = (torch.stack(allst).max(0).values - torch.stack(allst).min(0).values).max().item()
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
        # buckets launched from inside the backward pass (eagerly and from a replayed CUDA graph); the fp32
        # atomics of the weight-gradient kernels make runs differ in the last bits, hence the same tolerance
        assert o["stat_mean_err"] < 1e-6 and o["stat_spread"] > 1e-6, (o["stat_mean_err"], o["stat_spread"])
        for k in ("fused_overlap", "fused_overlap_graph", "nccl_central", "fused_central"):
            assert o[k][:2] == outs[0][k][:2], "replicas are not bit-identical"
            assert abs(o[k][0] - o["nccl"][0]) < 1e-4 * max(1.0, abs(o["nccl"][1])), (k, o[k], o["nccl"])
    print("multicast (NVLS) used by the fused path:", outs[0]["fused"][2])


FED_WORKER = r'''
import sys, json, torch
sys.path.insert(0, %r)
from idc_models_b200.parallel.world import World
from idc_models_b200.workloads import create_small_cnn
from idc_models_b200.fed.secure_fused import FusedSecureFederation
world = World.from_env()
dev = world.device
res = {}
# ---- secure federation: fused kernel vs library path vs float mean, percent = 1.0 and 0.5 -----------------
for percent in (1.0, 0.5):
    outs = {}
    for comm in ("fused", "nccl"):
        tpl = create_small_cnn(dev, seed=0, lr=1e-3)
        fed = FusedSecureFederation(world, 4 * world.size, percent, template=tpl, batch_size=16, seed=3, comm=comm)
        g = torch.Generator(device=dev).manual_seed(10 + world.rank)
        for j in range(fed.k_local):
            fed.set_client_data(j, torch.rand(48, 10, 10, 3, device=dev, generator=g),
                                (torch.rand(48, device=dev, generator=g) > 0.5).float())
        fed.client_fit(1)
        k, n = fed.k_local, fed.stride
        local = fed.params.tensor[:k * n].view(k, n).double().sum(0)
        world.all_reduce_sum(local)
        want = (local / fed.n_clients).float()
        got = fed.aggregate().clone()
        torch.cuda.synchronize()
        outs[comm] = ((got - want).abs().max().item(), got.double().sum().item())
        fed.client_update()
        assert torch.equal(fed.params.tensor[:n], fed.out.tensor[:n])
    res["secure_%%s" %% percent] = outs
# ---- FedAvg: fused kernel vs library aggregation on identical client blocks ---------------------------------
from idc_models_b200.fed import fedavg as F
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import RMSprop
from idc_models_b200.parallel.strategy import FederatedStrategy
strategy = FederatedStrategy(world)
sums = {}
for comm in ("fused", "nccl"):
    def model_fn():
        with strategy.scope():
            m = build_transfer_model("mobilenet_v2_tiny", (50, 50, 3), 1, device=dev, seed=0)
            m.base.trainable = True
        return F.from_keras_model(m, loss="bce", metrics=["accuracy"])
    fed = F.build_federated_averaging_process(model_fn, client_optimizer_fn=lambda: RMSprop(lr=1e-3), comm=comm)
    g = torch.Generator(device=dev).manual_seed(50 + world.rank)
    data = []
    for cid in range(2 * world.size):
        if cid %% world.size != world.rank:
            data.append([]); continue
        n_ex = 32 + 16 * (cid %% 3)          # unequal client weights
        x = torch.rand(n_ex, 50, 50, 3, device=dev, generator=g); y = (torch.rand(n_ex, device=dev, generator=g) > 0.5).float()
        data.append([(x[s:s + 16], y[s:s + 16]) for s in range(0, n_ex, 16)])
    state = fed.initialize()
    for _ in range(2):
        state, metrics = fed.next(state, data)
    torch.cuda.synchronize()
    sums[comm] = (state.master.double().sum().item(), state.master.double().abs().sum().item(), metrics[1])
res["fedavg"] = sums
allres = world.all_gather_object(res)
if world.rank == 0:
    print("RESULT " + json.dumps(allres))
world.barrier()
world.shutdown()
''' % str(ROOT)


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_federated_kernels_across_processes():
    """FedAvg weighted mean and the secure masked sum over NVLink peer memory: every rank receives the same
    model; fused == library path; secure mean == float mean within the fixed-point step."""
    script = ROOT / "build" / "_fed_worker.py"
    script.parent.mkdir(exist_ok=True)
    script.write_text(FED_WORKER)
    n = min(NGPU, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    outs = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    print(outs[0])
    for o in outs:
        for key in ("secure_1.0", "secure_0.5"):
            for comm in ("fused", "nccl"):
                assert o[key][comm][0] < 2.0 ** -15, (key, comm, o[key][comm])
                assert o[key][comm][1] == outs[0][key][comm][1], "ranks disagree on the aggregated model"
        f, l = o["fedavg"]["fused"], o["fedavg"]["nccl"]
        assert f[:2] == outs[0]["fedavg"]["fused"][:2], "ranks disagree on the FedAvg server model"
        assert abs(f[0] - l[0]) < 1e-3 * max(1.0, l[1]), (f, l)
