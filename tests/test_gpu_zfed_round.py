"""Single-GPU checks of what was added after this round's GPU budget was spent: the one-launch FedAvg round
(csrc/comm.cu:fedavg_round_kernel), the in-kernel mini-batch gather of the small CNN
(csrc/small_cnn.cu:idcb_small_cnn_step_indexed) and the copy-stream input stager of the fused executors
(models/fused/common.py:InputStager). Their first run on hardware is the driver's, so

* each check runs in its OWN python process (a device fault there cannot poison the CUDA context of the
  tests that follow), and
* the file sorts after every other GPU test file and before tests/test_zz_no_alloc.py.

Oracles: the PyTorch statement of the aggregation (fed/fedavg.py:weighted_delta_mean) and the dense launch of the
same kernel fed by an explicit ``torch.gather``. The multi-GPU version of the first check is part of
tests/test_gpu_multi.py::test_federated_kernels_across_processes.
"""
import json
import subprocess
import sys
from pathlib import Path

import pytest

# Not yet run on hardware when committed (see the module docstring). Non-strict xfail: a pass is reported as XPASS, a
# failure as XFAIL, and neither stops a `pytest -x` run before tests/test_zz_no_alloc.py, which must stay the last file
# collected and must get to run.
pytestmark = [pytest.mark.gpu,
              pytest.mark.xfail(strict=False, reason="written after the round's GPU budget was spent; the driver's run "
                                                     "is its first on a B200 (XPASS = it works)")]
ROOT = Path(__file__).resolve().parent.parent

PRELUDE = r'''
import sys, json, ctypes as C, torch
sys.path.insert(0, %r)
from idc_models_b200.ops import _lib
from idc_models_b200.parallel.world import World
world = World.from_env("cuda")
res = {}
''' % str(ROOT)

FEDAVG_ROUND = PRELUDE + r'''
from idc_models_b200.fed import fedavg
from idc_models_b200.parallel.symm import SymmAllocator
torch.manual_seed(2)
k_local, n = 4, 8192
alloc = SymmAllocator(world)
clients, weights = alloc.alloc(k_local * n, torch.float32), alloc.alloc(4, torch.float32)
glob, shadow = alloc.alloc(n, torch.float32), alloc.alloc(n, torch.bfloat16)
c = torch.randn(k_local, n, device="cuda")
c[1, 5000] = float("-inf")                     # dropped: inside the second range
c[2, 100] = float("nan")                       # kept: outside every aggregated range
wts = torch.tensor([3000., 1000., 2500., 500.], device="cuda")
clients.tensor.copy_(c.reshape(-1)); weights.tensor.copy_(wts)
base = torch.randn(n, device="cuda")
ranges = [(128, 2048), (4096, 6144), (8000, 8192)]
offs = (C.c_longlong * 3)(*[b for b, _ in ranges]); lens = (C.c_longlong * 3)(*[e - b for b, e in ranges])
cw = alloc.comm_world(use_barrier=True)        # a real grid barrier: the phases depend on each other across CTAs
lib = _lib.cuda_lib()
errs = []
for rep in range(3):                           # back to back: the pad epochs must chain
    out = torch.full((n,), float("nan"), device="cuda")
    _lib.check(lib.idcb_fedavg_round(C.byref(cw), clients.ptrs(), weights.ptrs(), glob.ptrs(), shadow.ptrs(), None, None,
                                     C.c_int(k_local), C.c_longlong(n), C.c_int(3), offs, lens, C.c_float(1.0),
                                     C.c_void_p(base.data_ptr()), C.c_void_p(out.data_ptr()), C.c_longlong(n), _lib.stream()))
    torch.cuda.synchronize()
    want = fedavg.weighted_delta_mean(base.cpu(), c.cpu(), wts.cpu(), 1.0, ranges)
    errs.append((out.cpu() - want).abs().max().item())
res["weights"] = weights.tensor[:4].tolist()
res["err"] = max(errs)
prev, kept = 0, True
for b, e in ranges + [(n, n)]:
    kept = kept and torch.equal(out[prev:b], base[prev:b])
    prev = e
res["outside_kept"] = kept
res["shadow_ok"] = all(torch.equal(shadow.tensor[b:e], out[b:e].to(torch.bfloat16)) for b, e in ranges)
bad = (C.c_longlong * 3)(128, 64, 8000)
rc = lib.idcb_fedavg_round(C.byref(cw), clients.ptrs(), weights.ptrs(), glob.ptrs(), shadow.ptrs(), None, None,
                           C.c_int(k_local), C.c_longlong(n), C.c_int(3), bad, lens, C.c_float(1.0),
                           C.c_void_p(base.data_ptr()), C.c_void_p(out.data_ptr()), C.c_longlong(n), _lib.stream())
res["bad_ranges_refused"] = rc != 0 and b"ascending" in lib.idcb_last_error()
print("RESULT " + json.dumps(res))
'''

INDEXED_CNN = PRELUDE + r'''
from idc_models_b200.fed.secure_fused import FusedSecureFederation
from idc_models_b200.workloads import create_small_cnn
tpl = create_small_cnn(world.device, seed=0, lr=1e-3)
k, n_ex, b = 4, 37, 16
fed = FusedSecureFederation(world, k, 1.0, template=tpl, batch_size=b, seed=9)
g = torch.Generator(device="cuda").manual_seed(11)
X = torch.rand(k, n_ex, 10, 10, 3, device="cuda", generator=g)
Y = (torch.rand(k, n_ex, device="cuda", generator=g) > 0.5).float()
perm = torch.stack([torch.randperm(n_ex, device="cuda", generator=g) for _ in range(k)])
out = []
for col0, bb in ((0, b), (16, b), (32, n_ex - 32)):
    fed.step += 1
    idx = perm[:, col0:col0 + bb]
    xb = torch.gather(X, 1, idx[:, :, None, None, None].expand(k, bb, 10, 10, 3)).contiguous()
    yb = torch.gather(Y, 1, idx).contiguous()
    fed.acc.zero_(); fed.grads.zero_()
    fed._launch(xb, yb, bb, True)
    torch.cuda.synchronize()
    ref = (fed._logits[:k * bb].clone(), fed.grads.clone(), fed.acc.clone())
    fed.acc.zero_(); fed.grads.zero_(); fed._logits.zero_()
    fed._launch_indexed(X, Y, perm, col0, n_ex, bb)
    torch.cuda.synchronize()
    out.append({"logits_equal": bool(torch.equal(fed._logits[:k * bb], ref[0])),
                "grad_scale": ref[1].abs().max().item(),
                "grad_err": (fed.grads - ref[1]).abs().max().item(),
                "counts_equal": bool(torch.equal(fed.acc[:, 1:], ref[2][:, 1:])),
                "loss_err": (fed.acc[:, 0] - ref[2][:, 0]).abs().max().item()})
res["steps"] = out
print("RESULT " + json.dumps(res))
'''


ASYNC_STAGE = PRELUDE + r'''
from idc_models_b200.models import applications as apps
from idc_models_b200.models.applications import build_transfer_model
from idc_models_b200.optim.optimizers import SGD
torch.manual_seed(0)
out = {}
for name in ("small_cnn", "mobilenet_v2_tiny"):
    if name == "small_cnn":
        m = apps.Sequential([apps.Conv2D(32, (3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                             apps.MaxPooling2D((2, 2)), apps.Dropout(0.0), apps.Flatten(),
                             apps.Dense(8, activation="relu"), apps.Dropout(0.0), apps.Dense(1)],
                            device="cuda", seed=3, executor="fused")
        shape = (32, 10, 10, 3)
    else:
        m = build_transfer_model(name, (50, 50, 3), 1, device="cuda", seed=0)
        shape = (16, 50, 50, 3)
    m.compile(optimizer=SGD(0.0), loss="bce", metrics=["accuracy"])
    ex = m.executor
    ex.use_graphs = name != "small_cnn"
    xs = [torch.rand(*shape).pin_memory() for _ in range(7)]
    y = (torch.rand(shape[0]) > 0.5).float().pin_memory()
    assert ex.async_stage
    if name == "small_cnn":
        run = lambda x, yy: ex.eval_batch(x, yy)[1]                      # eager inference (returns a copy)
    else:
        # the captured training step (learning rate 0, so every call sees the same weights); its BatchNorm sums are
        # accumulated atomically, hence the comparison below is relative to the batch-to-batch spread
        run = lambda x, yy: ex.train_batch(x, yy, shape[0])[1].clone()
    got = [run(x, y) for x in xs]                         # back to back, no synchronisation: both staging buffers cycle
    torch.cuda.synchronize()
    staged = hasattr(ex._get_plan(shape[0]), "_stager")
    want = [run(x.cuda(), y.cuda()) for x in xs]
    torch.cuda.synchronize()
    err = max((a - b).abs().max().item() for a, b in zip(got, want))
    spread = min((want[i] - want[i + 1]).abs().max().item() for i in range(6))   # different batches, different logits
    out[name] = {"staged": staged, "err": err, "spread": spread, "scale": max(w.abs().max().item() for w in want)}
res["models"] = out
print("RESULT " + json.dumps(res))
'''


def _run(name: str, body: str) -> dict:
    script = ROOT / "build" / name
    script.parent.mkdir(exist_ok=True)
    script.write_text(body)
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    return json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])


def test_fedavg_round_single_launch_single_gpu():
    """one launch = finite scan + weighted mean of every range + private copy of the whole model"""
    o = _run("_fedavg_round_worker.py", FEDAVG_ROUND)
    assert o["weights"] == [3000.0, 0.0, 2500.0, 500.0], o      # only the client with a bad value INSIDE a range is dropped
    assert o["err"] < 1e-5 and o["outside_kept"] and o["shadow_ok"] and o["bad_ranges_refused"], o


def test_small_cnn_indexed_launch_equals_gathered_launch():
    """the mini-batch gathered inside the kernel gives the logits (bit for bit), and the atomically accumulated
    gradients and loss sums (to rounding), of an explicit index-gather followed by the dense launch; same seed and
    step, so the dropout masks agree"""
    o = _run("_indexed_cnn_worker.py", INDEXED_CNN)
    for s in o["steps"]:
        assert s["logits_equal"] and s["counts_equal"], s
        assert s["grad_scale"] > 0 and s["grad_err"] <= 1e-5 * max(1.0, s["grad_scale"]) and s["loss_err"] < 1e-4, s


def test_pinned_host_input_through_the_copy_stream_stager():
    """a pinned host batch staged on the copy stream (double-buffered, moved into the graph's input buffer by a
    device-to-device copy) gives the logits of the same batch passed as a device tensor, for a run of different
    batches issued back to back, eagerly and under CUDA-graph replay"""
    o = _run("_async_stage_worker.py", ASYNC_STAGE)
    for name, r in o["models"].items():
        assert r["staged"], name
        assert r["spread"] > 0, (name, r)
        assert r["err"] <= 0.02 * r["spread"] + 1e-5 * max(1.0, r["scale"]), (name, r)
