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
# MobileNetV2: buckets announced per inverted-residual block from inside the backward pass
for comm in ("nccl", "fused_overlap", "fused_overlap_graph"):
    strategy = MirroredStrategy(world=world, comm=comm.split("_")[0])
    strategy.overlap = "overlap" in comm
    strategy.overlap_min_bucket = 2048
    with strategy.scope():
        m = build_transfer_model("mobilenet_v2_tiny", device=world.device, seed=0)
        m.base.trainable = True
        m.compile(optimizer=SGD(learning_rate=1e-2), loss="bce", metrics=["accuracy"])
    m.executor.use_graphs = comm.endswith("graph")
    if "overlap" in comm:
        assert strategy._overlap is not None and len(strategy._overlap.plan) >= 2, "no MobileNetV2 overlap schedule"
        res["mb_buckets"] = len(strategy._overlap.plan)
    g = torch.Generator().manual_seed(0)
    gb = 32 * world.size
    x = torch.rand(gb, 50, 50, 3, generator=g).to(world.device); y = (torch.rand(gb, generator=g) > 0.5).float().to(world.device)
    for _ in range(3):
        m.train_on_batch(x, y)
    torch.cuda.synchronize()
    res["mb_" + comm] = (m.store.master.double().sum().item(), m.store.master.double().abs().sum().item())
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
        for k in ("mb_fused_overlap", "mb_fused_overlap_graph"):
            assert o[k] == outs[0][k], "MobileNetV2 replicas are not bit-identical"
            assert abs(o[k][0] - o["mb_nccl"][0]) < 1e-4 * max(1.0, abs(o["mb_nccl"][1])), (k, o[k], o["mb_nccl"])
        assert o["mb_buckets"] >= 2
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
# ---- the one-launch FedAvg round: two ranges, unequal weights, a non-finite client on the last rank -----------
import ctypes as C
from idc_models_b200.ops import _lib
from idc_models_b200.parallel.symm import SymmAllocator
alloc = SymmAllocator(world, True)
W, R, kl, nn = world.size, world.rank, 3, 4096
cl, wt = alloc.alloc(kl * nn, torch.float32), alloc.alloc(4, torch.float32)
gb, gs = alloc.alloc(nn, torch.float32), alloc.alloc(nn, torch.bfloat16)
g = torch.Generator(device=dev).manual_seed(77)              # the same numbers on every rank
allc = torch.randn(W, kl, nn, device=dev, generator=g)
base = torch.randn(nn, device=dev, generator=g)
allc[W - 1, 1, 700] = float("nan")
allw = (1000.0 + 500.0 * torch.arange(W * kl, device=dev)).view(W, kl)
cl.tensor[:kl * nn].copy_(allc[R].reshape(-1)); wt.tensor[:kl].copy_(allw[R])
torch.cuda.synchronize(); world.barrier()
ranges = [(256, 1024), (2048, 3584)]
offs = (C.c_longlong * 2)(*[b for b, _ in ranges]); lens = (C.c_longlong * 2)(*[e - b for b, e in ranges])
cw = alloc.comm_world()
round_out = []
for rep in range(3):                                           # back to back: epochs and barriers must chain
    out = torch.full((nn,), float("nan"), device=dev)
    _lib.check(_lib.cuda_lib().idcb_fedavg_round(C.byref(cw), cl.ptrs(), wt.ptrs(), gb.ptrs(), gs.ptrs(), gb.mc(), gs.mc(),
        C.c_int(kl), C.c_longlong(nn), C.c_int(2), offs, lens, C.c_float(0.5), C.c_void_p(base.data_ptr()),
        C.c_void_p(out.data_ptr()), C.c_longlong(nn), _lib.stream()))
    round_out.append(out)
torch.cuda.synchronize()
want = F.weighted_delta_mean(base.cpu(), allc.view(-1, nn).cpu(), allw.view(-1).cpu(), 0.5, ranges)
res["fedavg_round"] = {"err": max((o.cpu() - want).abs().max().item() for o in round_out),
                       "outside_kept": all(torch.equal(o[:256], base[:256]) and torch.equal(o[1024:2048], base[1024:2048])
                                           and torch.equal(o[3584:], base[3584:]) for o in round_out),
                       "dropped_weight": wt.tensor[1].item(), "sum": round_out[-1].double().sum().item(),
                       "shadow_ok": all(torch.equal(gs.tensor[b:e], round_out[-1][b:e].to(torch.bfloat16)) for b, e in ranges)}
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
    for r, o in enumerate(outs):
        fr = o["fedavg_round"]
        assert fr["err"] < 1e-5 and fr["outside_kept"] and fr["shadow_ok"], fr
        assert fr["sum"] == outs[0]["fedavg_round"]["sum"], "ranks disagree on the one-launch FedAvg model"
        assert fr["dropped_weight"] == (0.0 if r == len(outs) - 1 else outs[r]["fedavg_round"]["dropped_weight"])
    assert outs[-1]["fedavg_round"]["dropped_weight"] == 0.0 and outs[0]["fedavg_round"]["dropped_weight"] > 0.0


STRESS_WORKER = r'''
import sys, json, ctypes as C, torch, torch.distributed as dist
sys.path.insert(0, %r)
from idc_models_b200.ops import _lib
from idc_models_b200.parallel.symm import SymmAllocator
from idc_models_b200.parallel.world import World
world = World.from_env()
dev, W, R = world.device, world.size, world.rank
lib = _lib.cuda_lib()
alloc = SymmAllocator(world)
NMAX = 1 << 16
grad, master, shadow = alloc.alloc(NMAX, torch.float32), alloc.alloc(NMAX, torch.float32), alloc.alloc(NMAX, torch.bfloat16)
s1 = torch.zeros(NMAX, device=dev)
ref = torch.zeros(NMAX, device=dev)             # what NCCL + a plain SGD step gives
bad = torch.zeros(4, dtype=torch.int64, device=dev)
gsh = torch.Generator().manual_seed(1234)        # shared by all ranks: the sequence of (n, offset, owner)
gme = torch.Generator(device=dev).manual_seed(77 + R)
mc = bool(grad.mc_ptr and master.mc_ptr and shadow.mc_ptr)
vp = C.c_void_p
ITERS = %d
special = [4, 8, 4 * W - 4, 4 * W + 4, 4 * (W + 1) * 3, 1020, 4]
for it in range(ITERS):
    if it < len(special):
        n = max(4, special[it])
    else:
        n = int(torch.randint(1, 4096, (1,), generator=gsh)) * 4
    b = int(torch.randint(0, (NMAX - n) // 4 + 1, (1,), generator=gsh)) * 4
    owner = int(torch.randint(0, W, (1,), generator=gsh)) if it %% 5 == 4 else -1
    # integer-valued gradients: every summation order gives the same fp32 bits, so the comparison is exact
    g = torch.randint(-8, 9, (n,), device=dev, generator=gme).float()
    grad.tensor[b:b + n].copy_(g)
    tot = g.clone()
    dist.all_reduce(tot)
    ref[b:b + n] += -0.5 * tot
    cw = alloc.comm_world()
    _lib.check(lib.idcb_allreduce_optim(
        C.byref(cw), grad.ptrs(b), master.ptrs(b), shadow.ptrs(b), grad.mc(b) if mc else vp(0),
        master.mc(b) if mc else vp(0), shadow.mc(b) if mc else vp(0), vp(s1.data_ptr() + 4 * b), vp(0),
        C.c_longlong(n), C.c_int(1), C.c_float(0.5), C.c_float(0.0), C.c_float(0.0), C.c_float(1e-7),
        C.c_float(1.0), C.c_int(owner), None, vp(0), _lib.stream()))
    m = master.tensor
    bad[0] += (m != ref).sum()                                     # element-wise, the whole buffer (untouched parts too)
    bad[1] += (grad.tensor != 0).sum()                             # the consumed bucket is zero again, nothing else was dirtied
    bad[2] += (shadow.tensor[b:b + n] != m[b:b + n].to(torch.bfloat16)).sum()
torch.cuda.synchronize()
chk = int(master.tensor.view(torch.int32).to(torch.int64).sum().item())
outs = world.all_gather_object((bad.tolist(), chk, float(ref.abs().sum().item())))
if R == 0:
    print("RESULT " + json.dumps({"outs": outs, "multicast": mc, "iters": ITERS}))
world.barrier()
world.shutdown()
'''


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_fused_allreduce_stress_randomised_ranges_elementwise_vs_nccl():
    """2,000 back-to-back launches of the fused all-reduce + optimizer kernel over random (offset, length, owner)
    ranges — including n = 4, lengths that are not a multiple of world x 4 and the reduce-to-owner variant —
    compared ELEMENT BY ELEMENT against NCCL all-reduce + the same SGD step. Gradients are small integers, so
    the comparison is bit-exact whatever order the switch / the ring adds them in. This is the race check of
    the epoch-flag barrier protocol (compute-sanitizer is closed on the pool)."""
    script = ROOT / "build" / "_stress_worker.py"
    script.parent.mkdir(exist_ok=True)
    script.write_text(STRESS_WORKER % (str(ROOT), 2000))
    n = min(NGPU, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    print(res)
    for bad, chk, mag in res["outs"]:
        assert bad[:3] == [0, 0, 0], f"mismatching elements (master, grad-not-zero, shadow): {bad}"
        assert chk == res["outs"][0][1], "replicas differ"
        assert mag > 0
