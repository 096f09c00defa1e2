"""GPU numerics of the fused communication kernels on ONE GPU: N ranks are emulated by one process
(N sets of buffers, the kernel is run once per emulated rank with the cross-rank barriers disabled — valid
because ranks own disjoint shards; see include/idcb200.h). Oracles are plain PyTorch / the host C++
restatement. The real multi-process NVLink path is exercised by tests/test_gpu_multi.py (needs >= 2 GPUs)."""
import ctypes as C

import numpy as np
import pytest
import torch

from idc_models_b200.fed import fedavg, secure
from idc_models_b200.ops import _lib
from idc_models_b200.parallel.fused_comm import allreduce_apply_emulated
from idc_models_b200.parallel.symm import EmulatedWorld

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,world", [("rmsprop", 4), ("sgd", 2), ("adam", 8), ("rmsprop", 3)])
def test_allreduce_optim_matches_sum_then_optimizer(kind, world):
    torch.manual_seed(0)
    n = 4096 * 3 + 64
    emu = EmulatedWorld(world)
    bufs = {"grad": emu.alloc(n, torch.float32), "master": emu.alloc(n, torch.float32),
            "shadow": emu.alloc(n, torch.bfloat16)}
    w0 = torch.randn(n, device="cuda")
    grads = [torch.randn(n, device="cuda") for _ in range(world)]
    for r in range(world):
        bufs["master"][r].tensor.copy_(w0)
        bufs["grad"][r].tensor.copy_(grads[r])
    states = [(torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda") if kind == "adam" else None)
              for _ in range(world)]
    lr, p1, p2, eps = {"rmsprop": (1e-2, 0.9, 0.0, 1e-7), "sgd": (1e-2, 0.0, 0.0, 0.0),
                       "adam": (1e-2 * (1 - 0.999) ** 0.5 / (1 - 0.9), 0.9, 0.999, 1e-7)}[kind]
    allreduce_apply_emulated(emu, bufs, states, n, kind, lr, p1, p2, eps, 0.25)
    torch.cuda.synchronize()
    g = sum(grads) * 0.25
    if kind == "rmsprop":
        want = w0 - lr * g / ((0.1 * g * g).sqrt() + eps)
    elif kind == "sgd":
        want = w0 - lr * g
    else:
        want = w0 - lr * (0.1 * g) / ((0.001 * g * g).sqrt() + eps)
    for r in range(world):
        assert torch.allclose(bufs["master"][r].tensor, want, rtol=1e-5, atol=1e-6), f"rank {r}"
        assert torch.equal(bufs["master"][r].tensor, bufs["master"][0].tensor), "replicas must be bit-identical"
        assert torch.equal(bufs["shadow"][r].tensor, bufs["master"][0].tensor.to(torch.bfloat16))
        assert bufs["grad"][r].tensor.abs().max().item() == 0.0, "the bucket must be re-zeroed"


def test_fedavg_reduce_matches_weighted_delta_mean():
    torch.manual_seed(1)
    world, k_local, n = 2, 3, 2048
    emu = EmulatedWorld(world)
    clients = emu.alloc(k_local * n, torch.float32)
    weights = emu.alloc(4, torch.float32)
    glob = emu.alloc(n, torch.float32)
    shadow = emu.alloc(n, torch.bfloat16)
    w0 = torch.randn(n, device="cuda")
    allc, allw = [], []
    for r in range(world):
        c = torch.randn(k_local, n, device="cuda")
        if r == 1:
            c[2, 300] = float("inf")            # this client must be dropped
        wts = torch.tensor([3000., 1000., 2500.], device="cuda") * (r + 1)
        clients[r].tensor.copy_(c.reshape(-1))
        weights[r].tensor[:k_local].copy_(wts)
        glob[r].tensor.copy_(w0)
        allc.append(c); allw.append(wts)
    lib = _lib.cuda_lib()
    b, e = 256, 1792                              # aggregate a sub-range (trainable suffix)
    for r in range(world):
        _lib.check(lib.idcb_fedavg_finite_flags(C.c_void_p(clients[r].tensor.data_ptr()), C.c_longlong(n), C.c_longlong(b),
                                                C.c_longlong(e - b), C.c_int(k_local),
                                                C.c_void_p(weights[r].tensor.data_ptr()), _lib.stream()))
    for r in range(world):
        cw = emu.comm_world(r)
        _lib.check(lib.idcb_fedavg_reduce(C.byref(cw), clients[r].ptrs(), weights[r].ptrs(), glob[r].ptrs(),
                                          shadow[r].ptrs(), None, None, C.c_int(k_local), C.c_longlong(n),
                                          C.c_longlong(b), C.c_longlong(e - b), C.c_float(1.0), _lib.stream()))
    torch.cuda.synchronize()
    assert weights[1].tensor[2].item() == 0.0
    want = fedavg.weighted_delta_mean(w0.cpu(), torch.cat(allc).cpu(), torch.cat(allw).cpu(), 1.0, [(b, e)])
    for r in range(world):
        assert torch.allclose(glob[r].tensor.cpu(), want, rtol=1e-4, atol=1e-5)
        assert torch.equal(glob[r].tensor[:b], w0[:b]) and torch.equal(glob[r].tensor[e:], w0[e:])


@pytest.mark.parametrize("world,k_local", [(2, 4), (8, 8), (1, 5)])
def test_secagg_masked_sum_matches_host_oracle(world, k_local):
    rng = np.random.RandomState(2)
    n, stride, off = 512, 1024, 256
    total = world * k_local
    w = (rng.randn(total, stride) * 2).astype(np.float32)
    emu = EmulatedWorld(world)
    clients = emu.alloc(k_local * stride, torch.float32)
    ybuf = emu.alloc(n, torch.int32)
    out = emu.alloc(stride, torch.float32)
    for r in range(world):
        clients[r].tensor.copy_(torch.from_numpy(w[r * k_local:(r + 1) * k_local].reshape(-1)).cuda())
    seed = 0x1234567890ABCDEF
    for r in range(world):     # phase A of every rank must precede phase B of any rank: run twice
        secure.fused_secure_mean(None, clients[r], ybuf[r], out[r], k_local, stride, off, n, seed, comm_world=emu.comm_world(r))
    for r in range(world):
        secure.fused_secure_mean(None, clients[r], ybuf[r], out[r], k_local, stride, off, n, seed, comm_world=emu.comm_world(r))
    torch.cuda.synchronize()
    # oracle: full masked vectors on the host -> per-GPU partial ring sums and the final mean
    y = secure.mask_vectors(w[:, off:off + n], seed)
    for r in range(world):
        part = secure.ring_sum(list(y[r * k_local:(r + 1) * k_local]))
        got = ybuf[r].tensor.cpu().numpy().view(np.uint32)[:n]
        assert np.array_equal(got, part), f"masked partial sum of GPU {r} differs from the host restatement"
    want = secure.dequantise_mean(secure.ring_sum(list(y)), total)
    for r in range(world):
        res = out[r].tensor.cpu().numpy()
        assert np.abs(res[off:off + n] - want).max() < 1e-6
        assert np.abs(res[off:off + n] - w[:, off:off + n].mean(0)).max() < 1e-4
        assert np.all(res[:off] == 0) and np.all(res[off + n:] == 0)
    if world > 1:
        # a single GPU's partial sum is still masked: it must not equal the plain quantised partial sum
        plain = np.rint(w[:k_local, off:off + n].astype(np.float64) * 65536).astype(np.int64).sum(0).astype(np.uint32)
        assert np.mean(ybuf[0].tensor.cpu().numpy().view(np.uint32)[:n] == plain) < 0.01


def test_fused_secure_federation_single_gpu():
    """8 resident clients train concurrently (one launch per step), aggregate through the masked ring sum and
    adopt the mean; percent selects the protected tensor prefix (secure_fed_model.py:115-121)."""
    from idc_models_b200.fed.secure_fused import FusedSecureFederation
    from idc_models_b200.parallel.world import World
    from idc_models_b200.workloads import create_small_cnn
    world = World.from_env("cuda")
    for percent, comm in ((1.0, "fused"), (0.5, "fused"), (0.0, "fused"), (1.0, "nccl")):
        tpl = create_small_cnn(world.device, seed=0, lr=1e-3)
        fed = FusedSecureFederation(world, 8, percent, template=tpl, batch_size=16, seed=1, comm=comm)
        assert fed.n_protected_tensors == int(6 * percent)
        g = torch.Generator(device="cuda").manual_seed(5)
        for j in range(8):
            fed.set_client_data(j, torch.rand(40, 10, 10, 3, device="cuda", generator=g),
                                (torch.rand(40, device="cuda", generator=g) > 0.5).float())
        before = fed.params.tensor[:8 * fed.stride].clone()
        stats = fed.client_fit(2)
        assert stats.shape == (8, 3) and torch.all(stats[:, 2] == 40)
        after = fed.params.tensor[:8 * fed.stride].view(8, fed.stride)
        assert not torch.equal(before.view(8, -1), after)            # every client moved ...
        assert not torch.equal(after[0], after[1])                   # ... along its own data
        want = after.double().mean(0).float()
        got = fed.aggregate().clone()
        torch.cuda.synchronize()
        assert (got - want).abs().max().item() < 2.0 ** -15, (percent, comm)
        fed.client_update()
        assert torch.equal(fed.params.tensor[:fed.stride], got)
        loss, acc = fed.evaluate(torch.rand(64, 10, 10, 3), (torch.rand(64) > 0.5).float(), steps=4)
        assert 0.0 < loss < 5.0 and 0.0 <= acc <= 1.0
