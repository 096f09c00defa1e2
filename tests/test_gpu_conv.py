"""GPU numerics of the tcgen05 implicit-GEMM convolutions against a plain PyTorch fp32 reference
of the same op, at the shapes the three model families use."""
import pytest
import torch

from idc_models_b200.ops import conv as C

pytestmark = pytest.mark.gpu


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _close(got, want, tol=2e-2):
    got = got.float()
    want = want.float()
    err = (got - want).abs().max().item()
    ref = want.abs().max().item() + 1e-6
    assert err <= tol * ref, f"max abs err {err:.4g} vs ref max {ref:.4g}"


FWD_CASES = [
    # b, h, w, cin, cout, k, stride, padding
    (2, 8, 8, 64, 64, 1, 1, "same"),
    (3, 13, 13, 256, 128, 1, 1, "same"),
    (2, 12, 12, 64, 64, 3, 1, "same"),
    (4, 13, 13, 128, 32, 3, 1, "same"),
    (2, 6, 6, 96, 128, 1, 1, "same"),
    (2, 25, 25, 64, 256, 3, 1, "same"),
    (5, 4, 4, 576, 160, 1, 1, "same"),
    (2, 7, 7, 24, 144, 1, 1, "same"),
    (2, 3, 3, 512, 512, 3, 1, "same"),
    (128, 1, 1, 992, 128, 1, 1, "same"),
]


@pytest.mark.parametrize("b,h,w,cin,cout,k,stride,padding", FWD_CASES)
def test_fprop_plain(b, h, w, cin, cout, k, stride, padding):
    torch.manual_seed(0)
    g = C.geom(h, w, cin, cout, k, stride, padding)
    x = _rand(b, h, w, cin)
    wt = _rand(cout, k, k, cin, scale=(k * k * cin) ** -0.5)
    y = torch.zeros(b, g.ho, g.wo, cout, device="cuda", dtype=torch.bfloat16)
    C.conv_fprop_raw(g, x, wt, y)
    torch.cuda.synchronize()
    _close(y, C.conv_ref(g, x.float(), wt.float()))


def test_fprop_fused_densenet_pattern():
    """BN+ReLU prologue on a channel prefix of a concat buffer, raw output written at a channel
    offset, batch statistics accumulated in the epilogue."""
    torch.manual_seed(1)
    b, h, w, ctot, cin, cout = 4, 13, 13, 256, 96, 128
    g = C.geom(h, w, cin, cout, 1)
    buf = _rand(b, h, w, ctot)
    wt = _rand(cout, 1, 1, cin, scale=cin ** -0.5)
    sc = torch.rand(cin, device="cuda") + 0.5
    sh = torch.randn(cin, device="cuda") * 0.3
    y = torch.zeros(b, h, w, cout, device="cuda", dtype=torch.bfloat16)
    s1 = torch.zeros(cout, device="cuda")
    s2 = torch.zeros(cout, device="cuda")
    C.conv_fprop_raw(g, buf, wt, y, pro_scale=sc, pro_shift=sh, pro_act=C.ACT_RELU, stat_sum=s1,
                     stat_sq=s2)
    torch.cuda.synchronize()
    a = torch.relu(buf[..., :cin].float() * sc + sh).to(torch.bfloat16).float()
    want = C.conv_ref(g, a, wt.float())
    _close(y, want)
    yf = y.float().reshape(-1, cout)
    _close(s1, yf.sum(0), 1e-3)
    _close(s2, (yf * yf).sum(0), 1e-3)
    # 3x3 128->32 written into the buffer at channel offset cin
    g3 = C.geom(h, w, cout, 32, 3)
    w3 = _rand(32, 3, 3, cout, scale=(9 * cout) ** -0.5)
    sc3 = torch.rand(cout, device="cuda") + 0.5
    sh3 = torch.randn(cout, device="cuda") * 0.3
    before = buf.clone()
    C.conv_fprop_raw(g3, y, w3, buf, y_coff=cin, pro_scale=sc3, pro_shift=sh3, pro_act=C.ACT_RELU)
    torch.cuda.synchronize()
    a3 = torch.relu(y.float() * sc3 + sh3).to(torch.bfloat16).float()
    _close(buf[..., cin:cin + 32], C.conv_ref(g3, a3, w3.float()))
    assert torch.equal(buf[..., :cin], before[..., :cin])
    assert torch.equal(buf[..., cin + 32:], before[..., cin + 32:])


def test_fprop_epilogue_bias_relu_residual():
    torch.manual_seed(2)
    b, h, w, cin, cout = 2, 12, 12, 64, 64
    g = C.geom(h, w, cin, cout, 3)
    x = _rand(b, h, w, cin)
    wt = _rand(cout, 3, 3, cin, scale=(9 * cin) ** -0.5)
    es = torch.rand(cout, device="cuda") + 0.5
    eh = torch.randn(cout, device="cuda")
    res = _rand(b, h, w, cout)
    y = torch.zeros(b, h, w, cout, device="cuda", dtype=torch.bfloat16)
    C.conv_fprop_raw(g, x, wt, y, epi_scale=es, epi_shift=eh, epi_act=C.ACT_RELU6, residual=res)
    torch.cuda.synchronize()
    want = torch.clamp(C.conv_ref(g, x.float(), wt.float()) * es + eh, 0, 6) + res.float()
    _close(y, want)


DGRAD_CASES = [
    (2, 12, 12, 64, 64, 3),
    (4, 13, 13, 128, 32, 3),
    (3, 13, 13, 256, 128, 1),
    (2, 6, 6, 512, 512, 3),
    (2, 7, 7, 24, 144, 1),
    (128, 1, 1, 992, 128, 1),
]


def _dgrad_ref(g, dy, wt, xshape):
    x = torch.zeros(xshape, device="cuda", requires_grad=True)
    y = C.conv_ref(g, x, wt.float())
    (gx,) = torch.autograd.grad(y, x, dy.float())
    return gx


@pytest.mark.parametrize("b,h,w,cin,cout,k", DGRAD_CASES)
def test_dgrad_plain(b, h, w, cin, cout, k):
    torch.manual_seed(3)
    g = C.geom(h, w, cin, cout, k)
    wt = _rand(cout, k, k, cin, scale=(k * k * cout) ** -0.5)
    dy = _rand(b, g.ho, g.wo, cout)
    dx = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    C.conv_dgrad_raw(g, dy, wt, dx)
    torch.cuda.synchronize()
    _close(dx, _dgrad_ref(g, dy, wt, (b, h, w, cin)))


def test_dgrad_relu_mask_and_bias_grad():
    torch.manual_seed(4)
    b, h, w, cin, cout = 2, 12, 12, 64, 128
    g = C.geom(h, w, cin, cout, 3)
    wt = _rand(cout, 3, 3, cin, scale=(9 * cout) ** -0.5)
    dy = _rand(b, h, w, cout)
    xin = torch.relu(_rand(b, h, w, cin))
    dx = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    db = torch.zeros(cin, device="cuda")
    C.conv_dgrad_raw(g, dy, wt, dx, mode=1, aux=xin, stat_a=db)
    torch.cuda.synchronize()
    want = _dgrad_ref(g, dy, wt, (b, h, w, cin)) * (xin.float() > 0)
    _close(dx, want)
    _close(db, dx.float().reshape(-1, cin).sum(0), 5e-3)


def test_dgrad_bn_relu_backward_into_g():
    """mode 2: BN+ReLU backward fused in the dgrad epilogue, accumulated into the dense block's
    fp32 gradient buffer at a channel prefix."""
    torch.manual_seed(5)
    b, h, w, ctot, cin, cout = 4, 13, 13, 256, 96, 128
    g = C.geom(h, w, cin, cout, 1)
    wt = _rand(cout, 1, 1, cin, scale=cout ** -0.5)
    dy = _rand(b, h, w, cout)
    xraw = _rand(b, h, w, ctot)
    sc = torch.rand(cin, device="cuda") + 0.5
    sh = torch.randn(cin, device="cuda") * 0.3
    G = torch.randn(b, h, w, ctot, device="cuda")
    G0 = G.clone()
    dpre = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    sa = torch.zeros(cin, device="cuda")
    sb = torch.zeros(cin, device="cuda")
    C.conv_dgrad_raw(g, dy, wt, dpre, mode=2, aux=xraw, bwd_scale=sc, bwd_shift=sh,
                     bwd_act=C.ACT_RELU, gacc=G, stat_a=sa, stat_b=sb)
    torch.cuda.synchronize()
    da = _dgrad_ref(g, dy, wt, (b, h, w, cin))
    xr = xraw[..., :cin].float()
    want = da * ((xr * sc + sh) > 0)
    _close(dpre, want)
    _close(G[..., :cin] - G0[..., :cin], want * sc, 2e-2)
    assert torch.equal(G[..., cin:], G0[..., cin:])
    _close(sa, want.reshape(-1, cin).sum(0), 2e-2)
    _close(sb, (want * xr).reshape(-1, cin).sum(0), 2e-2)


WGRAD_CASES = [
    (2, 8, 8, 64, 64, 1),
    (4, 13, 13, 128, 32, 3),
    (3, 13, 13, 256, 128, 1),
    (2, 12, 12, 64, 64, 3),
    (2, 6, 6, 96, 128, 1),
    (2, 7, 7, 24, 144, 1),
    (2, 3, 3, 512, 512, 3),
    (16, 25, 25, 64, 128, 3),
]


@pytest.mark.parametrize("b,h,w,cin,cout,k", WGRAD_CASES)
def test_wgrad(b, h, w, cin, cout, k):
    torch.manual_seed(6)
    g = C.geom(h, w, cin, cout, k)
    x = _rand(b, h, w, cin)
    dy = _rand(b, g.ho, g.wo, cout, scale=0.1)
    dw = torch.zeros(cout, k, k, cin, device="cuda")
    C.conv_wgrad_raw(g, x, dy, dw)
    torch.cuda.synchronize()
    wref = torch.zeros(cout, k, k, cin, device="cuda", requires_grad=True)
    y = C.conv_ref(g, x.float(), wref)
    (gw,) = torch.autograd.grad(y, wref, dy.float())
    _close(dw, gw, 1e-2)


def test_wgrad_prologue_and_offsets():
    torch.manual_seed(7)
    b, h, w, ctot, cin, cout = 4, 13, 13, 256, 96, 32
    g = C.geom(h, w, cin, cout, 3)
    buf = _rand(b, h, w, ctot)
    dybuf = _rand(b, h, w, 64, scale=0.1)
    sc = torch.rand(cin, device="cuda") + 0.5
    sh = torch.randn(cin, device="cuda") * 0.3
    dw = torch.zeros(cout, 3, 3, cin, device="cuda")
    C.conv_wgrad_raw(g, buf, dybuf, dw, dy_coff=32, pro_scale=sc, pro_shift=sh, pro_act=C.ACT_RELU)
    torch.cuda.synchronize()
    a = torch.relu(buf[..., :cin].float() * sc + sh).to(torch.bfloat16).float()
    wref = torch.zeros(cout, 3, 3, cin, device="cuda", requires_grad=True)
    y = C.conv_ref(g, a, wref)
    (gw,) = torch.autograd.grad(y, wref, dybuf[..., 32:64].float())
    _close(dw, gw, 1e-2)


# ---------------------------------------------------------------------------------------------
# Shapes that only the whole-model tests used to reach (round-1 review): strided / asymmetric stems with a
# tap-packed K axis, wgrad with Cout that is not a multiple of 64, tiny-M mode-2 dgrad.
# ---------------------------------------------------------------------------------------------
STEM_CASES = [
    # b, h, w, cin, cout, k, stride, padding
    (4, 50, 50, 8, 64, 7, 2, ((3, 3), (3, 3))),      # DenseNet stem on IDC patches (RGB padded to 8)
    (4, 32, 32, 8, 64, 7, 2, ((3, 3), (3, 3))),      # ... on CIFAR
    (4, 50, 50, 8, 32, 3, 2, ((0, 1), (0, 1))),      # MobileNetV2 stem, Keras correct_pad on an even size
    (4, 25, 25, 8, 32, 3, 2, ((1, 1), (1, 1))),      # ... odd size
    (2, 13, 13, 96, 96, 3, 2, ((0, 1), (0, 1))),
]


@pytest.mark.parametrize("b,h,w,cin,cout,k,stride,padding", STEM_CASES)
def test_fprop_strided_and_tap_packed(b, h, w, cin, cout, k, stride, padding):
    torch.manual_seed(10)
    g = C.geom(h, w, cin, cout, k, stride, padding)
    x = _rand(b, h, w, cin)
    wt = _rand(cout, k, k, cin, scale=(k * k * cin) ** -0.5)
    y = torch.zeros(b, g.ho, g.wo, cout, device="cuda", dtype=torch.bfloat16)
    s1 = torch.zeros(cout, device="cuda")
    C.conv_fprop_raw(g, x, wt, y, stat_sum=s1)
    torch.cuda.synchronize()
    _close(y, C.conv_ref(g, x.float(), wt.float()))
    _close(s1, y.float().reshape(-1, cout).sum(0), 1e-3)
    # and its weight gradient (the same linearised (tap, channel) axis, now as the M axis)
    dy = _rand(b, g.ho, g.wo, cout, scale=0.1)
    dw = torch.zeros(cout, k, k, cin, device="cuda")
    C.conv_wgrad_raw(g, x, dy, dw)
    torch.cuda.synchronize()
    wref = torch.zeros(cout, k, k, cin, device="cuda", requires_grad=True)
    (gw,) = torch.autograd.grad(C.conv_ref(g, x.float(), wref), wref, dy.float())
    _close(dw, gw, 1e-2)


@pytest.mark.parametrize("cout", [24, 96, 160])
def test_wgrad_cout_not_a_multiple_of_64(cout):
    torch.manual_seed(11)
    b, h, w, cin = 4, 7, 7, 192
    g = C.geom(h, w, cin, cout, 1)
    x = _rand(b, h, w, cin)
    dy = _rand(b, h, w, cout, scale=0.1)
    dw = torch.zeros(cout, 1, 1, cin, device="cuda")
    C.conv_wgrad_raw(g, x, dy, dw)
    torch.cuda.synchronize()
    wref = torch.zeros(cout, 1, 1, cin, device="cuda", requires_grad=True)
    (gw,) = torch.autograd.grad(C.conv_ref(g, x.float(), wref), wref, dy.float())
    _close(dw, gw, 1e-2)


def _mode2_reference(g, dy, wt, xraw, cin, sc, sh):
    da = _dgrad_ref(g, dy, wt, (xraw.shape[0], g.h, g.w, cin))
    xr = xraw[..., :cin].float()
    return da * ((xr * sc + sh) > 0), xr


@pytest.mark.parametrize("b,h,w", [(128, 1, 1), (1, 1, 1), (14, 3, 3)])
def test_dgrad_mode2_tiny_m(b, h, w):
    """M = 128 (exactly one tile), M = 1 and M = 126 (a tile with invalid rows)."""
    torch.manual_seed(12)
    cin, cout = 128, 32
    g = C.geom(h, w, cin, cout, 3)
    wt = _rand(cout, 3, 3, cin, scale=(9 * cout) ** -0.5)
    dy = _rand(b, h, w, cout)
    xraw = _rand(b, h, w, cin)
    sc = torch.rand(cin, device="cuda") + 0.5
    sh = torch.randn(cin, device="cuda") * 0.3
    dpre = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    sa = torch.zeros(cin, device="cuda")
    sb = torch.zeros(cin, device="cuda")
    C.conv_dgrad_raw(g, dy, wt, dpre, mode=2, aux=xraw, bwd_scale=sc, bwd_shift=sh, bwd_act=C.ACT_RELU,
                     stat_a=sa, stat_b=sb)
    torch.cuda.synchronize()
    want, xr = _mode2_reference(g, dy, wt, xraw, cin, sc, sh)
    _close(dpre, want)
    _close(sa, want.reshape(-1, cin).sum(0), 2e-2)
    _close(sb, (want * xr).reshape(-1, cin).sum(0), 2e-2)


def _bn_backward_reference(dpre, x, mean, rstd, scale):
    """dx, dgamma, dbeta of y = scale * (x - mean) * rstd / ... with batch statistics (rows = dim 0)"""
    n = dpre.shape[0]
    xhat = (x - mean) * rstd
    dbeta = dpre.sum(0)
    dgamma = (dpre * xhat).sum(0)
    dx = scale * (dpre - dbeta / n - xhat * dgamma / n)
    return dx, dgamma, dbeta


@pytest.mark.parametrize("b,h,w,fused", [(4, 13, 13, True), (128, 1, 1, True), (14, 3, 3, True),
                                         (128, 13, 13, False)])
def test_dgrad_mode2_with_batchnorm_tail(b, h, w, fused):
    """The BatchNorm backward finished inside the dgrad launch (grid barrier, tile parked in TMEM) equals
    dgrad followed by the closed-form BN backward. The last case has more tiles than SMs: the native side must
    take its two-launch route through dx and still produce the same numbers."""
    torch.manual_seed(13)
    cin, cout = 128, 32
    g = C.geom(h, w, cin, cout, 3)
    wt = _rand(cout, 3, 3, cin, scale=(9 * cout) ** -0.5)
    dy = _rand(b, h, w, cout)
    y1 = _rand(b, h, w, cin)
    yf = y1.float().reshape(-1, cin)
    mean = yf.mean(0)
    rstd = (yf.var(0, unbiased=False) + 1e-5).rsqrt()
    gamma = torch.rand(cin, device="cuda") + 0.5
    beta = torch.randn(cin, device="cuda") * 0.3
    sc = gamma * rstd
    sh = beta - mean * sc
    dpre = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(b, h, w, cin, device="cuda", dtype=torch.bfloat16)
    sa = torch.zeros(cin, device="cuda")
    sb = torch.zeros(cin, device="cuda")
    dgamma = torch.full((cin,), 0.25, device="cuda")       # accumulated into, not overwritten
    dbeta = torch.full((cin,), -0.5, device="cuda")
    counter = torch.zeros(64, device="cuda")
    for rep in range(2):        # a second launch with a fresh counter word gives the same answer
        sa.zero_(); sb.zero_(); counter.zero_()
        tail = C.bn_tail(out, mean, rstd, b * h * w, counter=counter, dgamma=dgamma, dbeta=dbeta)
        C.conv_dgrad_raw(g, dy, wt, dpre, mode=2, aux=y1, bwd_scale=sc, bwd_shift=sh, bwd_act=C.ACT_RELU,
                         stat_a=sa, stat_b=sb, tail=tail)
    torch.cuda.synchronize()
    want_dpre, xr = _mode2_reference(g, dy, wt, y1, cin, sc, sh)
    dx, dg, db = _bn_backward_reference(want_dpre.reshape(-1, cin), xr.reshape(-1, cin), mean, rstd, sc)
    _close(out.reshape(-1, cin), dx, 3e-2)
    _close(dgamma - 0.25, 2 * dg, 3e-2)
    _close(dbeta + 0.5, 2 * db, 3e-2)
    if fused:
        assert dpre.abs().max().item() == 0, "dpre must not be written when the tail runs in the kernel"
    # frozen BatchNorm: dx = scale * dpre, no statistics terms
    out2 = torch.zeros_like(out)
    tail = C.bn_tail(out2, mean, rstd, b * h * w, frozen=True)
    C.conv_dgrad_raw(g, dy, wt, dpre, mode=2, aux=y1, bwd_scale=sc, bwd_shift=sh, bwd_act=C.ACT_RELU,
                     stat_a=sa, stat_b=sb, tail=tail)
    torch.cuda.synchronize()
    _close(out2, want_dpre * sc, 2e-2)


def test_centre_tap_window_equals_the_full_3x3_on_one_pixel():
    """A 3x3 SAME convolution over 1x1 images through the weight-window descriptor (1x1 over w[:, 1, 1, :]):
    forward, data gradient and weight gradient equal the full 3x3 reference; the other taps' gradients stay zero."""
    torch.manual_seed(14)
    b, cin, cout = 128, 128, 32
    g3 = C.geom(1, 1, cin, cout, 3)
    g = C.centre_tap(g3)
    x = _rand(b, 1, 1, cin)
    wt = _rand(cout, 3, 3, cin, scale=cin ** -0.5)
    y = torch.zeros(b, 1, 1, cout, device="cuda", dtype=torch.bfloat16)
    sc = torch.rand(cin, device="cuda") + 0.5
    sh = torch.randn(cin, device="cuda") * 0.3
    C.conv_fprop_raw(g, x, wt, y, pro_scale=sc, pro_shift=sh, pro_act=C.ACT_RELU)
    a = torch.relu(x.float() * sc + sh).to(torch.bfloat16).float()
    want = C.conv_ref(g3, a, wt.float())
    torch.cuda.synchronize()
    _close(y, want)
    dy = _rand(b, 1, 1, cout)
    dx = torch.zeros(b, 1, 1, cin, device="cuda", dtype=torch.bfloat16)
    C.conv_dgrad_raw(g, dy, wt, dx)
    torch.cuda.synchronize()
    _close(dx, _dgrad_ref(g3, dy, wt, (b, 1, 1, cin)))
    dw = torch.zeros(cout, 3, 3, cin, device="cuda")
    C.conv_wgrad_raw(g, x, dy, dw, pro_scale=sc, pro_shift=sh, pro_act=C.ACT_RELU)
    torch.cuda.synchronize()
    wref = torch.zeros(cout, 3, 3, cin, device="cuda", requires_grad=True)
    (gw,) = torch.autograd.grad(C.conv_ref(g3, a, wref), wref, dy.float())
    _close(dw, gw, 1e-2)
    off = dw.clone()
    off[:, 1, 1, :] = 0
    assert off.abs().max().item() == 0


@pytest.mark.parametrize("b,h,w,cin", [(128, 3, 3, 512), (128, 1, 1, 992), (32, 6, 6, 96)])
def test_dgrad_mode2_with_slice_tail(b, h, w, cin):
    """The 1x1 dgrad that is the last writer of a DenseNet block-buffer slice also emits that slice's finished
    gradient and its BatchNorm's bookkeeping: equal to dgrad + idcb_bn_bwd_slice with the pending fold."""
    from idc_models_b200.ops import kernels as K
    torch.manual_seed(15)
    ctot, cout = 1024, 128
    M = b * h * w
    g = C.geom(h, w, cin, cout, 1)
    wt = _rand(cout, 1, 1, cin, scale=cout ** -0.5)
    dy = _rand(b, h, w, cout)
    buf = _rand(b, h, w, ctot)
    bf_ = buf.float().reshape(-1, ctot)
    mean = bf_.mean(0)
    rstd = (bf_.var(0, unbiased=False) + 1e-5).rsqrt()
    gamma = torch.rand(cin, device="cuda") + 0.5
    sc = gamma * rstd[:cin]
    sh = torch.randn(cin, device="cuda") * 0.1 - mean[:cin] * sc
    G0 = torch.randn(b, h, w, ctot, device="cuda")
    s1_0, s2_0 = torch.randn(ctot, device="cuda"), torch.randn(ctot, device="cuda")
    c0 = cin - 32

    def run(fused):
        G = G0.clone()
        s1, s2 = s1_0.clone(), s2_0.clone()
        sa, sb = torch.zeros(cin, device="cuda"), torch.zeros(cin, device="cuda")
        dgm, dbt = torch.ones(cin, device="cuda"), torch.ones(cin, device="cuda")
        out = torch.zeros(b, h, w, 32, device="cuda", dtype=torch.bfloat16)
        ctr = torch.zeros(64, device="cuda")
        st = None
        if fused:
            assert C.slice_tail_fits(M, cin)
            st = C.slice_tail(out, c0, s1, s2, mean, rstd, mean[:cin].contiguous(), rstd[:cin].contiguous(), M, ctr,
                              dgamma=dgm, dbeta=dbt)
        C.conv_dgrad_raw(g, dy, wt, None, dx_pitch=ctot, mode=2, aux=buf, bwd_scale=sc, bwd_shift=sh,
                         bwd_act=C.ACT_RELU, gacc=G, stat_a=sa, stat_b=sb, slice_tail=st)
        if not fused:
            K.bn_bwd_slice(G, buf, out, c0, 32, float(M), s1, s2, mean, rstd,
                           pending=(sa, sb, mean[:cin].contiguous(), rstd[:cin].contiguous(), sc, dgm, dbt))
        torch.cuda.synchronize()
        return G, s1, s2, dgm, dbt, out

    Gf, s1f, s2f, dgf, dbf, outf = run(True)
    Gr, s1r, s2r, dgr, dbr, outr = run(False)
    _close(Gf, Gr, 1e-3)
    _close(outf, outr, 2e-2)
    _close(dgf, dgr, 1e-3)
    _close(dbf, dbr, 1e-3)
    outside = torch.ones(ctot, dtype=torch.bool, device="cuda")
    outside[c0:c0 + 32] = False
    _close(s1f[outside], s1r[outside], 1e-3)
    _close(s2f[outside], s2r[outside], 1e-3)
