"""Integration on CPU, 2 replicas over gloo: north-star config #1 (MobileNetV2, synthetic IDC patches, global
batch 32): replicas stay bit-identical and the 2-replica loss equals the single-replica loss on the same
global batch (SURVEY.md section 4 "Integration, CPU 2 replicas")."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent

WORKER = r'''
import os, sys, json, torch
sys.path.insert(0, %r)
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import RMSprop
from idc_models_b200.parallel.strategy import MirroredStrategy, CentralStorageStrategy
from idc_models_b200.parallel.world import World
kind = sys.argv[1]
world = World.from_env("cpu")
strategy = (CentralStorageStrategy if kind == "central" else MirroredStrategy)(world=world)
with strategy.scope():
    m = build_transfer_model("mobilenet_v2", device="cpu", seed=world.rank)   # different seeds: sync must fix it
    # frozen base (phase 1 of the transfer scripts): BatchNorm runs on moving statistics, so the result
    # must not depend on how the global batch is split (trainable BN uses per-replica batch statistics
    # by design, like MirroredStrategy)
    m.compile(optimizer=RMSprop(lr=1e-3), loss="bce", metrics=["accuracy"])
g = torch.Generator().manual_seed(0)
x = torch.rand(32, 50, 50, 3, generator=g); y = (torch.rand(32, generator=g) > 0.5).float()
h = m.fit([(x, y)] * 2, epochs=1, verbose=0)
chk = float(m.store.master.double().sum())
outs = world.all_gather_object((chk, h.history["loss"][0], strategy.num_replicas_in_sync))
if world.rank == 0:
    print("RESULT " + json.dumps(outs))
world.shutdown()
''' % str(ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(kind, nproc):
    script = ROOT / "build" / "_dist_worker.py"
    script.parent.mkdir(exist_ok=True)
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script), kind]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    import json
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


@pytest.mark.parametrize("kind", ["mirrored", "central"])
def test_two_cpu_replicas_stay_identical_and_match_single_replica(kind):
    two = _run(kind, 2)
    one = _run(kind, 1)
    assert two[0][2] == 2 and one[0][2] == 1
    assert two[0][0] == two[1][0], "replica weight checksums differ"
    assert abs(two[0][1] - one[0][1]) < 5e-4, f"2-replica loss {two[0][1]} vs single-replica {one[0][1]}"
    assert abs(two[0][0] - one[0][0]) < 1e-3 * max(1.0, abs(one[0][0]))


def test_overlap_bucket_plan_is_static_and_covers_the_range():
    """gradient buckets launched from inside the backward pass (parallel/fused_comm.py): contiguous, top-down,
    merged up to the minimum size, identical for every rank"""
    from idc_models_b200.parallel.fused_comm import plan_buckets
    offs = [900, 700, 650, 300, 100]                      # announced as the backward walks down
    plan = plan_buckets(offs, 0, 1000, min_bucket=200)
    assert plan == [(700, 700, 1000), (300, 300, 700), (100, 100, 300), (0, 0, 100)]
    # covers [lo, hi) exactly once, in descending order
    assert plan[0][2] == 1000 and plan[-1][1] == 0
    assert all(a[1] == b[2] for a, b in zip(plan, plan[1:]))
    # a trainable range that starts above 0 (frozen lower layers) and unaligned announcements
    plan = plan_buckets([801, 503, 90], 400, 1000, min_bucket=100)
    assert plan == [(800, 800, 1000), (500, 500, 800), (400, 400, 500)]
    assert plan_buckets([], 0, 64, 16) == [(0, 0, 64)]
    # executors expose the offsets statically
    from idc_models_b200.models.applications import build_transfer_model
    m = build_transfer_model("densenet_tiny", (50, 50, 3), 1, device="cpu", seed=0, executor="reference")
    assert not hasattr(m.executor, "grad_ready_offsets") or isinstance(m.executor.grad_ready_offsets(), list)


def test_jsonl_logger_and_device_timer(tmp_path):
    import json
    from idc_models_b200.utils.trace import DeviceTimer, JsonlLogger, nvtx_range
    path = tmp_path / "log" / "run.jsonl"
    log = JsonlLogger(str(path))
    with nvtx_range("unit"):
        with DeviceTimer() as t:
            sum(range(1000))
    log.log("epoch", epoch=0, loss=0.5)
    log.log("epoch", epoch=1, loss=0.25, ms=t.ms)
    recs = [json.loads(l) for l in path.read_text().splitlines()]
    assert [r["epoch"] for r in recs] == [0, 1] and recs[1]["loss"] == 0.25 and recs[1]["ms"] >= 0
    JsonlLogger(None).log("epoch", epoch=2)               # disabled logger: no file, no error
    JsonlLogger(str(path), rank=1).log("epoch", epoch=3)  # only rank 0 writes
    assert len(path.read_text().splitlines()) == 2


FED_WORKER = r'''
import sys, json, torch
sys.path.insert(0, %r)
from idc_models_b200.fed import fedavg as F
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import SGD
from idc_models_b200.parallel.strategy import FederatedStrategy
from idc_models_b200.parallel.world import World
placement = sys.argv[1]
world = World.from_env("cpu")
strategy = FederatedStrategy(world)
def model_fn():
    with strategy.scope():
        m = build_transfer_model("mobilenet_v2_tiny", device="cpu", seed=0)
        m.base.trainable = False          # head only: deterministic, cheap
    return F.from_keras_model(m, loss="bce", metrics=["accuracy"])
g = torch.Generator().manual_seed(0)
def client(n_batches):
    return [(torch.rand(8, 50, 50, 3, generator=g), (torch.rand(8, generator=g) > 0.5).float()) for _ in range(n_batches)]
data = [client(4), client(1), client(1)]          # unequal clients: balanced puts the big one alone on a rank
fed = F.build_federated_averaging_process(model_fn, client_optimizer_fn=lambda: SGD(learning_rate=0.1), placement=placement)
state = fed.initialize()
state, metrics = fed.next(state, data)
ev = F.build_federated_evaluation(model_fn, placement=placement)(state.model, data)
mine = F._local_clients(data, world, placement)
outs = world.all_gather_object((float(state.master.double().sum()), metrics[1], ev[1], [len(d) for d in mine]))
if world.rank == 0:
    print("RESULT " + json.dumps(outs))
world.shutdown()
''' % str(ROOT)


def _run_fed(placement, nproc):
    import json
    script = ROOT / "build" / "_fed_cpu_worker.py"
    script.parent.mkdir(exist_ok=True)
    script.write_text(FED_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script), placement]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="2"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])


def test_fedavg_client_placement_does_not_change_the_result():
    """3 unequal clients on 2 gloo ranks: round-robin and balanced (LPT) placement aggregate to the same server model
    and the same example-weighted metrics as the single-process run; balanced gives the big client a rank of its own."""
    one = _run_fed("round_robin", 1)[0]
    rr = _run_fed("round_robin", 2)
    bal = _run_fed("balanced", 2)
    assert sorted(map(tuple, (o[3] for o in bal))) == [(1, 1), (4,)]
    assert sorted(map(tuple, (o[3] for o in rr))) == [(1,), (4, 1)]
    for outs in (rr, bal):
        for o in outs:
            assert abs(o[0] - one[0]) < 1e-4 * max(1.0, abs(one[0]))
            assert abs(o[1] - one[1]) < 1e-5 and abs(o[2] - one[2]) < 1e-5
