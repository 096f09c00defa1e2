"""GPU numerics of the non-GEMM kernels (BN bookkeeping, pooling, depthwise, head, optimizers) against
plain PyTorch fp32 statements of the same ops."""
import math

import pytest
import torch
import torch.nn.functional as F

from idc_models_b200.ops import conv as OC
from idc_models_b200.ops import kernels as K

pytestmark = pytest.mark.gpu
DEV = "cuda"


def bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(torch.bfloat16)


def close(a, b, tol=2e-2, what=""):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item()
    ref = b.abs().max().item() + 1e-6
    assert err <= tol * ref, f"{what} max err {err:.4g} vs ref max {ref:.4g}"


def test_bn_finalize_fold_and_apply():
    torch.manual_seed(0)
    c, m = 96, 500
    x = bf(m, c)
    xf = x.float()
    s, q = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    K.channel_stats(x.view(1, 1, m, c), c, s, q)
    close(s, xf.sum(0), 1e-3)
    close(q, (xf * xf).sum(0), 1e-3)
    gamma, beta = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV)
    mm, mv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    scale, shift = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    mean, rstd = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    K.bn_finalize(s, q, float(m), gamma, beta, 1e-3, scale, shift, mean, rstd, mm, mv, 0.9)
    var = xf.var(0, unbiased=False)
    close(mean, xf.mean(0), 1e-3)
    close(rstd, (var + 1e-3).rsqrt(), 1e-3)
    close(mm, 0.1 * xf.mean(0), 1e-3)
    close(mv, 0.9 + 0.1 * xf.var(0, unbiased=True), 1e-3)
    y = torch.empty_like(x)
    K.affine_act(x.view(1, 1, m, c), y.view(1, 1, m, c), c, scale, shift, 1)
    want = torch.relu(F.batch_norm(xf, None, None, gamma, beta, True, 0.0, 1e-3))
    close(y, want)
    K.bn_fold(gamma, beta, mm, mv, 1e-3, scale, shift)
    close(scale, gamma * (mv + 1e-3).rsqrt(), 1e-4)
    close(shift, beta - mm * gamma * (mv + 1e-3).rsqrt(), 1e-4)


def test_bn_backward_coeffs_and_affine2():
    """standalone BN+ReLU backward assembled from the pieces == autograd"""
    torch.manual_seed(1)
    c, m = 64, 777
    x = bf(m, c)
    dy = bf(m, c, scale=0.1)
    gamma, beta = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV) * 0.2
    xr = x.float().requires_grad_(True)
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    out = torch.relu(F.batch_norm(xr, None, None, g_, b_, True, 0.0, 1e-5))
    out.backward(dy.float())
    # our pieces
    s, q = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    K.channel_stats(x.view(1, 1, m, c), c, s, q)
    scale, shift = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    mean, rstd = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    K.bn_finalize(s, q, float(m), gamma, beta, 1e-5, scale, shift, mean, rstd)
    pre = x.float() * scale + shift
    dpre = (dy.float() * (pre > 0)).to(torch.bfloat16)
    sa, sb = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    K.channel_stats2(dpre.view(1, 1, m, c), x.view(1, 1, m, c), c, sa, sb)
    close(sa, dpre.float().sum(0), 1e-3)
    close(sb, (dpre.float() * x.float()).sum(0), 1e-3)
    dg, db = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    A, B, C0 = (torch.empty(c, device=DEV) for _ in range(3))
    K.bn_bwd_coeffs(sa, sb, mean, rstd, scale, float(m), dgamma=dg, dbeta=db, A=A, B=B, C0=C0)
    close(dg, g_.grad, 2e-2, "dgamma")
    close(db, b_.grad, 2e-2, "dbeta")
    dx = torch.empty_like(x)
    K.affine2(dpre.view(1, 1, m, c), x.view(1, 1, m, c), dx.view(1, 1, m, c), c, A, B, C0)
    close(dx, xr.grad, 3e-2, "dx")


@pytest.mark.parametrize("k,s,pad,is_max,h", [(2, 2, 0, True, 25), (2, 2, 0, True, 12), (3, 2, 1, True, 25),
                                              (2, 2, 0, False, 13), (2, 2, 0, False, 6)])
def test_pool_fwd_bwd(k, s, pad, is_max, h):
    torch.manual_seed(2)
    b, c = 3, 64
    x = torch.relu(bf(b, h, h, c)) if is_max else bf(b, h, h, c)
    ho = (h + 2 * pad - k) // s + 1
    y = torch.zeros(b, ho, ho, c, device=DEV, dtype=torch.bfloat16)
    K.pool_fwd(x, y, c, k, s, pad, is_max)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    if is_max:
        yr = F.max_pool2d(F.pad(xr, (pad, pad, pad, pad)), k, s)
    else:
        yr = F.avg_pool2d(xr, k, s)
    close(y, yr.permute(0, 2, 3, 1))
    dy = bf(b, ho, ho, c)
    dx = torch.zeros_like(x)
    sa = torch.zeros(c, device=DEV)
    K.pool_bwd(x, y, dy, dx, c, k, s, pad, is_max, relu_mask=is_max, stat_a=sa)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    want = xr.grad.permute(0, 2, 3, 1)
    if is_max:
        want = want * (x.float() > 0)
        # ties between equal bf16 values route the gradient to every tied element in our kernel;
        # compare only where the window maximum is unique
    got = dx.float()
    frac_bad = ((got - want).abs() > 2e-2 * (want.abs().max() + 1e-6)).float().mean().item()
    assert frac_bad < (0.02 if is_max else 1e-6), f"{frac_bad} of elements differ"
    close(sa, got.reshape(-1, c).sum(0), 1e-3)


def test_pool_max_with_bn_relu_prologue():
    torch.manual_seed(3)
    b, h, c = 2, 25, 64
    x = bf(b, h, h, c)
    sc, sh = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV) * 0.3
    y = torch.zeros(b, 13, 13, c, device=DEV, dtype=torch.bfloat16)
    K.pool_fwd(x, y, c, 3, 2, 1, True, scale=sc, shift=sh, act=1)
    a = torch.relu(x.float() * sc + sh).to(torch.bfloat16).float()
    want = F.max_pool2d(F.pad(a.permute(0, 3, 1, 2), (1, 1, 1, 1)), 3, 2).permute(0, 2, 3, 1)
    close(y, want, 1e-6)


@pytest.mark.parametrize("stride,h,c", [(1, 25, 32), (2, 25, 96), (1, 13, 144), (2, 13, 144), (1, 4, 576), (2, 4, 576), (1, 2, 960)])
def test_depthwise(stride, h, c):
    from idc_models_b200.models.spec import correct_pad
    torch.manual_seed(4)
    b = 3
    pad = "same" if stride == 1 else (correct_pad(h), correct_pad(h))
    g = OC.geom(h, h, c, c, 3, stride, pad)
    x = bf(b, h, h, c)
    w = bf(3, 3, c, scale=0.3)
    sc, sh = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV) * 0.3
    y = torch.zeros(b, g.ho, g.wo, c, device=DEV, dtype=torch.bfloat16)
    s1, s2 = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    K.dwconv_fprop(g, x, w, y, scale=sc, shift=sh, act=2, stat_sum=s1, stat_sq=s2)
    xr = x.float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    pre = xr * sc + sh
    a = torch.clamp(pre, 0, 6)
    an = F.pad(a.permute(0, 3, 1, 2), (g.pad_l, g.pad_r, g.pad_t, g.pad_b))
    yr = F.conv2d(an, wr.permute(2, 0, 1).unsqueeze(1), stride=stride, groups=c).permute(0, 2, 3, 1)
    close(y, yr)
    close(s1, y.float().reshape(-1, c).sum(0), 1e-3)
    close(s2, (y.float() ** 2).reshape(-1, c).sum(0), 1e-3)
    dy = bf(b, g.ho, g.wo, c, scale=0.1)
    yr.backward(dy.float())
    dx = torch.zeros_like(x)
    sa, sb = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    K.dwconv_dgrad(g, dy, w, x, dx, scale=sc, shift=sh, act=2, stat_a=sa, stat_b=sb)
    want_dpre = xr.grad / sc  # autograd gives d/dx = dpre * scale
    close(dx, want_dpre, 3e-2, "dpre")
    close(sa, dx.float().reshape(-1, c).sum(0), 1e-3)
    close(sb, (dx.float() * x.float()).reshape(-1, c).sum(0), 2e-3)
    dw = torch.zeros(3, 3, c, device=DEV)
    K.dwconv_wgrad(g, x, dy, dw, scale=sc, shift=sh, act=2)
    close(dw, wr.grad, 2e-2, "dw")


@pytest.mark.parametrize("k,loss_kind,hw,c", [(1, 0, 1, 1024), (1, 0, 4, 1280), (10, 1, 1, 1920), (1, 0, 1, 512)])
def test_head_forward_backward(k, loss_kind, hw, c):
    torch.manual_seed(5)
    b = 16
    x = bf(b, hw, 1, c)
    sc, sh = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV) * 0.3
    w = torch.randn(k, c, device=DEV) * c ** -0.5
    bias = torch.randn(k, device=DEV) * 0.1
    labels = (torch.rand(b, device=DEV) > 0.5).float() if loss_kind == 0 else torch.randint(0, k, (b,), device=DEV).float()
    logits = torch.zeros(b, k, device=DEV)
    acc = torch.zeros(4, device=DEV)
    dw, db = torch.zeros(k, c, device=DEV), torch.zeros(k, device=DEV)
    dx = torch.zeros(b, hw, 1, c, device=DEV, dtype=torch.bfloat16)
    G = torch.zeros(b, hw, 1, c, device=DEV)
    sa, sb = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    gs = 1.0 / b
    K.head(x, hw, c, k, w, bias, labels, logits, acc, pro_scale=sc, pro_shift=sh, pro_act=1,
           loss_kind=loss_kind, training=True, grad_scale=gs, dw=dw, db=db, dx=dx, gacc=G, stat_a=sa, stat_b=sb)
    xr = x.float().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), bias.clone().requires_grad_(True)
    pre = xr * sc + sh
    feat = torch.relu(pre).mean(dim=(1, 2))
    lg = feat @ wr.t() + br
    if loss_kind == 0:
        loss = F.binary_cross_entropy_with_logits(lg.reshape(-1), labels, reduction="sum")
        correct = ((lg.reshape(-1) > 0.5).float() == labels).sum()
    else:
        loss = F.cross_entropy(lg, labels.long(), reduction="sum")
        correct = (lg.argmax(-1) == labels.long()).sum()
    (loss * gs).backward()
    close(logits, lg, 1e-3, "logits")
    assert abs(acc[0].item() - loss.item()) < 1e-2 * max(1.0, loss.item())
    assert acc[1].item() == correct.item() and acc[2].item() == b
    close(dw, wr.grad, 1e-3, "dw")
    close(db, br.grad, 1e-3, "db")
    want_dpre = xr.grad / sc
    close(dx, want_dpre, 2e-2, "dpre")
    close(G, xr.grad, 2e-2, "G")
    close(sa, want_dpre.reshape(-1, c).sum(0), 2e-2, "sa")


@pytest.mark.parametrize("kind", ["rmsprop", "sgd", "adam"])
def test_optimizer_kernels_match_closed_form(kind):
    from idc_models_b200.models.applications import Sequential, Conv2D, Flatten, Dense, MaxPooling2D
    from idc_models_b200.optim import optimizers as O
    torch.manual_seed(6)
    n = 4096 + 64
    w0 = torch.randn(n)
    res = {}
    for dev in ("cpu", "cuda"):
        class S:
            pass
        st = S()
        st.master = w0.clone().to(dev)
        st.grad = torch.zeros(n, device=dev)
        st.shadow = torch.zeros(n, device=dev, dtype=torch.bfloat16) if dev == "cuda" else None
        opt = {"rmsprop": O.RMSprop(1e-2), "sgd": O.SGD(1e-2, momentum=0.9), "adam": O.Adam(1e-2)}[kind]
        opt.bind(st, [(0, 1024), (2048, n)])
        g = torch.Generator().manual_seed(7)
        for _ in range(3):
            st.grad.copy_(torch.randn(n, generator=g).to(dev))
            opt.apply(grad_scale=0.5, zero_grad=True)
        res[dev] = st
    close(res["cuda"].master.cpu(), res["cpu"].master, 1e-5, kind)
    assert torch.equal(res["cuda"].master[1024:2048].cpu(), w0[1024:2048])
    assert res["cuda"].grad[:1024].abs().max().item() == 0.0
    close(res["cuda"].shadow[:1024], res["cuda"].master[:1024], 1e-2)


def test_pack_input():
    x = torch.rand(4, 50, 50, 3, device=DEV)
    y = torch.ones(4, 50, 50, 8, device=DEV, dtype=torch.bfloat16)
    K.pack_input(x, y)
    assert torch.equal(y[..., :3], x.to(torch.bfloat16))
    assert y[..., 3:].abs().max().item() == 0.0


def test_small_cnn_fused_step_matches_reference_without_dropout():
    """whole-network kernel (conv, pool, dense x2, BCE, backward) vs the PyTorch fp32 executor"""
    from idc_models_b200.models import applications as apps
    from idc_models_b200.optim.optimizers import SGD
    nets = []
    for ex in ("fused", "reference"):
        m = apps.Sequential([apps.Conv2D(32, (3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                             apps.MaxPooling2D((2, 2)), apps.Dropout(0.0), apps.Flatten(),
                             apps.Dense(8, activation="relu"), apps.Dropout(0.0), apps.Dense(1)],
                            device="cuda", seed=3, executor=ex)
        m.compile(optimizer=SGD(0.0), loss="bce", metrics=["accuracy"])
        nets.append(m)
    f, r = nets
    r.store.master.copy_(f.store.master)
    torch.manual_seed(0)
    x = torch.rand(32, 10, 10, 3, device=DEV)
    y = (torch.rand(32, device=DEV) > 0.5).float()
    sf, lf = f.executor.train_batch(x, y, 32)
    sr, lr = r.executor.train_batch(x, y, 32)
    torch.cuda.synchronize()
    close(lf, lr, 1e-4, "logits")
    assert abs(sf[0].item() - sr[0].item()) < 1e-3 and sf[1].item() == sr[1].item() and sf[2].item() == 32
    for name in f.store.order:
        close(f.store.grad_view(name), r.store.grad_view(name), 1e-3, name)
    # evaluation path + dropout training run
    se, _ = f.executor.eval_batch(x, y)
    assert abs(se[0].item() - sr[0].item()) < 1e-3
    m = apps.Sequential([apps.Conv2D(32, (3, 3), activation="relu", input_shape=(10, 10, 3), strides=2),
                         apps.MaxPooling2D((2, 2)), apps.Dropout(0.25), apps.Flatten(),
                         apps.Dense(8, activation="relu"), apps.Dropout(0.5), apps.Dense(1)], device="cuda")
    m.compile(optimizer="rmsprop", loss="bce", metrics=["accuracy"])
    x[y > 0.5] *= 0.5
    h = m.fit([(x, y)] * 20, epochs=3, verbose=0)
    assert h.history["loss"][-1] < h.history["loss"][0]


def test_synth_patches_kernel():
    import ctypes as C
    from idc_models_b200.ops import _lib
    lib = _lib.cuda_lib()
    x = torch.zeros(64, 50, 50, 3, device=DEV)
    lab = torch.full((64,), -1.0, device=DEV)
    _lib.check(lib.idcb_synth_patches(C.c_void_p(x.data_ptr()), C.c_void_p(lab.data_ptr()), C.c_longlong(64), C.c_int(50),
                                      C.c_int(50), C.c_int(3), C.c_ulonglong(7), C.c_uint(1), _lib.stream()))
    torch.cuda.synchronize()
    assert 0.0 <= x.min().item() and x.max().item() <= 1.0
    assert set(lab.tolist()) <= {0.0, 1.0} and 10 < lab.sum().item() < 54
    pos, neg = x[lab > 0.5].mean().item(), x[lab < 0.5].mean().item()
    assert neg - pos > 0.05          # positives are darker: the classes are separable
    x2 = torch.zeros_like(x)
    _lib.check(lib.idcb_synth_patches(C.c_void_p(x2.data_ptr()), None, C.c_longlong(64), C.c_int(50), C.c_int(50),
                                      C.c_int(3), C.c_ulonglong(7), C.c_uint(1), _lib.stream()))
    assert torch.equal(x, x2)        # counter-based: reproducible


def test_bn_bwd_slice_folds_the_pending_consumer_batchnorm():
    """idcb_bn_bwd_slice == bn_bwd_coeffs (S1/S2/dgamma/dbeta of the pending consumer) followed by the mode-2
    slice backward, in one launch."""
    from idc_models_b200.ops import kernels as K
    torch.manual_seed(21)
    M, ctot, pc, c_lo, cn = 4 * 13 * 13, 256, 160, 128, 32
    G = torch.randn(4, 13, 13, ctot, device="cuda")
    x = (torch.randn(4, 13, 13, ctot, device="cuda")).to(torch.bfloat16)
    mean = torch.randn(ctot, device="cuda") * 0.1
    rstd = torch.rand(ctot, device="cuda") + 0.5
    p_sa, p_sb = torch.randn(pc, device="cuda"), torch.randn(pc, device="cuda")
    p_scale = torch.rand(pc, device="cuda") + 0.5
    s1_0, s2_0 = torch.randn(ctot, device="cuda"), torch.randn(ctot, device="cuda")
    # two-launch reference
    s1, s2 = s1_0.clone(), s2_0.clone()
    dg_ref, db_ref = torch.ones(pc, device="cuda"), torch.ones(pc, device="cuda")
    K.bn_bwd_coeffs(p_sa, p_sb, mean[:pc].contiguous(), rstd[:pc].contiguous(), p_scale, float(M), dgamma=dg_ref,
                    dbeta=db_ref, acc_s1=s1[:pc], acc_s2=s2[:pc])
    want = torch.zeros(4, 13, 13, cn, device="cuda", dtype=torch.bfloat16)
    K.bn_bwd_apply(2, G, x, want, cn, float(M), sa=s1[c_lo:c_lo + cn], sb=s2[c_lo:c_lo + cn],
                   mean=mean[c_lo:c_lo + cn], rstd=rstd[c_lo:c_lo + cn], g_coff=c_lo, x_coff=c_lo)
    # fused
    t1, t2 = s1_0.clone(), s2_0.clone()
    dg, db = torch.ones(pc, device="cuda"), torch.ones(pc, device="cuda")
    got = torch.zeros_like(want)
    K.bn_bwd_slice(G, x, got, c_lo, cn, float(M), t1, t2, mean, rstd,
                   pending=(p_sa, p_sb, mean[:pc].contiguous(), rstd[:pc].contiguous(), p_scale, dg, db))
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    assert torch.allclose(dg, dg_ref) and torch.allclose(db, db_ref)
    outside = torch.ones(ctot, dtype=torch.bool, device="cuda")
    outside[c_lo:c_lo + cn] = False
    assert torch.allclose(t1[outside], s1[outside]) and torch.allclose(t2[outside], s2[outside])
    # nothing pending: plain mode 2
    got2 = torch.zeros_like(want)
    K.bn_bwd_slice(G, x, got2, c_lo, cn, float(M), s1, s2, mean, rstd)
    torch.cuda.synchronize()
    assert torch.equal(got2, want)


def test_pool_max_backward_through_recorded_argmax():
    """Overlapping 3x3/2 max pool with the BatchNorm+ReLU prologue (DenseNet's pool1): the backward pass driven by
    the argmax bytes the forward pass recorded gives exactly what the value-comparing backward pass gives (first
    maximum wins a tie in both), including the BatchNorm-backward column sums."""
    torch.manual_seed(31)
    b, h, c = 4, 25, 64
    x = bf(b, h, h, c)
    # plenty of exact ties: quantise the input coarsely
    x = (x.float() * 2).round().div(2).to(torch.bfloat16)
    sc, sh = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV) * 0.3
    y = torch.zeros(b, 13, 13, c, device=DEV, dtype=torch.bfloat16)
    arg = torch.full((b, 13, 13, c), 255, device=DEV, dtype=torch.uint8)
    K.pool_fwd(x, y, c, 3, 2, 1, True, scale=sc, shift=sh, act=1, argmax=arg)
    y2 = torch.zeros_like(y)
    K.pool_fwd(x, y2, c, 3, 2, 1, True, scale=sc, shift=sh, act=1)
    assert torch.equal(y, y2) and int(arg.max()) <= 8
    dy = bf(b, 13, 13, c)
    outs = []
    for use_arg in (True, False):
        dx = torch.zeros(b, h, h, c, device=DEV, dtype=torch.bfloat16)
        sa, sb = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
        K.pool_bwd(x, y, dy, dx, c, 3, 2, 1, True, scale=sc, shift=sh, act=1, stat_a=sa, stat_b=sb,
                   argmax=arg if use_arg else None)
        torch.cuda.synchronize()
        outs.append((dx, sa, sb))
    assert torch.equal(outs[0][0], outs[1][0])
    close(outs[0][1], outs[1][1], 1e-4)
    close(outs[0][2], outs[1][2], 1e-4)
    # and against autograd where the window maximum is unique
    a = torch.relu(x.float() * sc + sh).to(torch.bfloat16).float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(F.pad(a, (1, 1, 1, 1)), 3, 2)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    want = a.grad.permute(0, 2, 3, 1) * ((x.float() * sc + sh) > 0)
    frac_bad = ((outs[0][0].float() - want).abs() > 2e-2 * (want.abs().max() + 1e-6)).float().mean().item()
    assert frac_bad < 0.02, frac_bad
