"""Whole-model GPU checks: the fused sm_100a pipelines (bf16 compute) against the plain PyTorch fp32
reference executor running the same weights on the same inputs — loss, logits and every gradient."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _pair(arch, input_shape=(50, 50, 3), units=1, fine_tune_at=0, seed=0):
    from idc_models_b200.models.applications import build_transfer_model
    from idc_models_b200.optim.optimizers import SGD
    nets = []
    for ex in ("fused", "reference_bf16"):
        net = build_transfer_model(arch, input_shape, units, device="cuda", seed=seed, executor=ex)
        net.base.trainable = True
        for l in net.base.layers[:fine_tune_at]:
            l.trainable = False
        nets.append(net)
    f, r = nets
    # identical, bf16-representable weights on both sides; non-trivial BN parameters and statistics
    g = torch.Generator(device="cpu").manual_seed(seed + 1)
    for name in f.store.order:
        s = f.store.slots[name]
        v = f.store.view(name)
        if s.kind == "gamma":
            v.copy_((0.5 + torch.rand(s.shape, generator=g)).to(v.device))
        elif s.kind in ("beta", "moving_mean", "bias"):
            v.copy_((0.2 * torch.randn(s.shape, generator=g)).to(v.device))
        elif s.kind == "moving_var":
            v.copy_((0.5 + torch.rand(s.shape, generator=g)).to(v.device))
    f.store.master.copy_(f.store.master.to(torch.bfloat16).float())
    f.store.refresh_shadow()
    r.store.master.copy_(f.store.master)
    r.store.stats.copy_(f.store.stats)
    for net in nets:
        net.compile(optimizer=SGD(lr=0.0), loss="bce" if units == 1 else "cce", metrics=["accuracy"])
    return f, r


def _compare(f, r, batch=32, units=1, tol_logit=0.08, min_cos=0.97):
    torch.manual_seed(3)
    h, w, c = f.input_shape
    x = torch.rand(batch, h, w, c, device="cuda")
    x = x.to(torch.bfloat16).float()
    y = (torch.rand(batch, device="cuda") > 0.5).float() if units == 1 else torch.randint(0, units, (batch,), device="cuda").float()
    sf, lf = f.executor.train_batch(x, y, batch)
    sr, lr = r.executor.train_batch(x, y if units == 1 else y.long(), batch)
    torch.cuda.synchronize()
    lf, lr = lf.float(), lr.float()
    scale = lr.abs().max().item() + 1e-3
    err = (lf - lr).abs().max().item()
    print(f"logit max err {err:.4g} (scale {scale:.4g}); loss fused {sf[0].item() / batch:.5f} ref {sr[0].item() / batch:.5f}")
    assert err <= tol_logit * max(scale, 1.0), f"logits differ: {err} vs scale {scale}"
    assert abs(sf[0].item() - sr[0].item()) / batch < 0.05
    bad = []
    worst = 1.0
    for name in f.store.order:
        s = f.store.slots[name]
        if not s.trainable_kind:
            continue
        if f.store.slots[name].layer not in f.trainable_layer_indices():
            continue
        gf = f.store.grad_view(name).float().reshape(-1)
        gr = r.store.grad_view(name).float().reshape(-1)
        nr = gr.norm().item()
        nf = gf.norm().item()
        if nr < 1e-7 and nf < 1e-6:
            continue
        cos = torch.dot(gf, gr).item() / (nf * nr + 1e-20)
        ratio = nf / (nr + 1e-20)
        worst = min(worst, cos)
        if cos < min_cos or not (0.8 < ratio < 1.25):
            bad.append((name, round(cos, 4), round(ratio, 4), nr))
    print("worst gradient cosine:", worst, "bad:", bad[:12])
    assert not bad, f"{len(bad)} gradient tensors disagree, first: {bad[:8]}"
    # BN moving statistics were updated identically
    if f.store.n_stats:
        ms_err = (f.store.stats - r.store.stats).abs().max().item()
        assert ms_err < 0.05, f"moving statistics differ by {ms_err}"


def test_vgg16_full_backward():
    f, r = _pair("vgg16")
    _compare(f, r)


def test_vgg16_fine_tune_at_15():
    f, r = _pair("vgg16", fine_tune_at=15)
    assert sum(t.numel() for t in f.trainable_weights) == 7079937
    _compare(f, r)


def test_densenet121_full_backward():
    f, r = _pair("densenet121")
    _compare(f, r)


def test_densenet121_fine_tune_at_150():
    f, r = _pair("densenet121", fine_tune_at=150)
    _compare(f, r)


def test_densenet201_cifar_shape_10_classes():
    f, r = _pair("densenet201", (32, 32, 3), units=10, fine_tune_at=150)
    _compare(f, r, units=10)


def test_fused_eval_matches_reference():
    f, r = _pair("densenet121")
    x = torch.rand(32, 50, 50, 3, device="cuda").to(torch.bfloat16).float()
    y = (torch.rand(32, device="cuda") > 0.5).float()
    sf, lf = f.executor.eval_batch(x, y)
    sr, lr = r.executor.eval_batch(x, y)
    assert (lf - lr).abs().max().item() < 0.08 * max(1.0, lr.abs().max().item())


def test_training_reduces_loss_densenet():
    from idc_models_b200.models.applications import build_transfer_model
    from idc_models_b200.optim.optimizers import RMSprop
    net = build_transfer_model("densenet121", device="cuda")
    net.base.trainable = True
    net.compile(optimizer=RMSprop(lr=1e-4), loss="bce", metrics=["accuracy"])
    torch.manual_seed(0)
    x = torch.rand(32, 50, 50, 3, device="cuda")
    y = (torch.rand(32, device="cuda") > 0.5).float()
    x[y > 0.5] *= 0.7
    h = net.fit([(x, y)] * 4, epochs=3, verbose=0)
    print(h.history)
    assert h.history["loss"][-1] < h.history["loss"][0]


def test_mobilenet_v2_full_backward():
    f, r = _pair("mobilenet_v2")
    _compare(f, r)


def test_mobilenet_v2_fine_tune_at_100():
    f, r = _pair("mobilenet_v2", fine_tune_at=100)
    assert sum(t.numel() for t in f.trainable_weights) == 1863873
    _compare(f, r)


def test_head_only_phase1_all_archs():
    """phase 1 of the transfer scripts: frozen base, only the Dense head trains"""
    for arch in ("vgg16", "densenet121", "mobilenet_v2"):
        f, r = _pair(arch, fine_tune_at=10 ** 6)
        _compare(f, r)
