"""Convolution front-end: geometry helpers, the raw bindings to the tcgen05 implicit-GEMM kernels
(``csrc/umma_conv.cu``) and the plain-PyTorch fp32 statement of the same math (CPU path + oracle).

Activation convention everywhere in this package: NHWC, i.e. a 4-D tensor ``[B, H, W, pitch]``
whose last dimension may be wider than the logical channel count (``coff``/``channels`` select a
slice). Conv weights are ``[Cout, KH, KW, Cin]``.

Reference parity: these are the cuDNN calls TensorFlow issues for ``tf.keras`` ``Conv2D`` inside
``tf.keras.applications`` models (/root/reference/dist_model_tf_vgg.py:119,
dist_model_tf_dense.py:131, dist_model_tf_mobile.py:119).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch
import torch.nn.functional as F

from . import _lib

ACT_NONE, ACT_RELU, ACT_RELU6 = 0, 1, 2


@dataclass(frozen=True)
class ConvGeom:
    """Forward geometry of one convolution (Keras semantics resolved to explicit padding)."""
    h: int
    w: int
    cin: int
    cout: int
    kh: int
    kw: int
    stride: int
    pad_t: int
    pad_l: int
    pad_b: int
    pad_r: int
    w_pitch: int = 0     # weight window (include/idcb200.h: idcb_conv_desc): row pitch / first element, 0 = dense
    w_off: int = 0

    @property
    def ho(self) -> int:
        return (self.h + self.pad_t + self.pad_b - self.kh) // self.stride + 1

    @property
    def wo(self) -> int:
        return (self.w + self.pad_l + self.pad_r - self.kw) // self.stride + 1


def same_padding(size: int, k: int, stride: int) -> tuple[int, int]:
    """TensorFlow 'SAME' padding (before, after) for one spatial dim."""
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return total // 2, total - total // 2


def geom(h: int, w: int, cin: int, cout: int, k: int, stride: int = 1, padding="same") -> ConvGeom:
    """padding: 'same' | 'valid' | ((top, bottom), (left, right)) explicit (ZeroPadding2D + VALID)."""
    if padding == "same":
        (pt, pb), (pl, pr) = same_padding(h, k, stride), same_padding(w, k, stride)
    elif padding == "valid":
        pt = pb = pl = pr = 0
    else:
        (pt, pb), (pl, pr) = padding
    return ConvGeom(h, w, cin, cout, k, k, stride, pt, pl, pb, pr)


def _desc(g: ConvGeom, batch: int, x_pitch: int, x_coff: int, y_pitch: int, y_coff: int) -> _lib.ConvDesc:
    return _lib.ConvDesc(batch, g.h, g.w, g.cin, g.ho, g.wo, g.cout, g.kh, g.kw, g.stride, g.pad_t,
                         g.pad_l, x_pitch, x_coff, y_pitch, y_coff, g.w_pitch, g.w_off)


def centre_tap(g: ConvGeom) -> ConvGeom:
    """A k x k stride-1 SAME convolution whose input is a single pixel only ever multiplies its centre tap (every
    other tap sees zero padding): the same result — forward, data gradient and weight gradient, the untouched
    taps' gradients being exactly zero — comes from a 1x1 convolution over the window w[:, k//2, k//2, :] of the
    weight tensor, with 1/k^2 of the MMA work."""
    assert g.h == 1 and g.w == 1 and g.stride == 1 and g.kh == g.kw and g.pad_t == g.kh // 2 == g.pad_l
    c = (g.kh // 2) * g.kw + g.kw // 2
    return ConvGeom(1, 1, g.cin, g.cout, 1, 1, 1, 0, 0, 0, 0, w_pitch=g.kh * g.kw * g.cin, w_off=c * g.cin)


def _chk(t: torch.Tensor, dtype, name: str) -> None:
    if t is None:
        return
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous CUDA {dtype} tensor, got {t.dtype} "
                         f"{t.device} contiguous={t.is_contiguous()}")


def conv_fprop_raw(g: ConvGeom, x: torch.Tensor, w: torch.Tensor, y: torch.Tensor, *, x_coff: int = 0,
                   y_coff: int = 0, pro_scale=None, pro_shift=None, pro_act: int = ACT_NONE,
                   epi_scale=None, epi_shift=None, epi_act: int = ACT_NONE, residual=None,
                   res_coff: int = 0, stat_sum=None, stat_sq=None, bn=None) -> None:
    """y[..., y_coff:y_coff+cout] = epi(conv(pro(x[..., x_coff:x_coff+cin]), w)) on the GPU."""
    _chk(x, torch.bfloat16, "x"); _chk(w, torch.bfloat16, "w"); _chk(y, torch.bfloat16, "y")
    for t, n in ((pro_scale, "pro_scale"), (pro_shift, "pro_shift"), (epi_scale, "epi_scale"),
                 (epi_shift, "epi_shift"), (stat_sum, "stat_sum"), (stat_sq, "stat_sq")):
        _chk(t, torch.float32, n)
    _chk(residual, torch.bfloat16, "residual")
    b = x.shape[0]
    d = _desc(g, b, x.shape[-1], x_coff, y.shape[-1], y_coff)
    f = _lib.ConvFused(_lib.ptr(pro_scale), _lib.ptr(pro_shift), pro_act, _lib.ptr(epi_scale),
                       _lib.ptr(epi_shift), epi_act, _lib.ptr(residual),
                       residual.shape[-1] if residual is not None else 0, res_coff,
                       _lib.ptr(stat_sum), _lib.ptr(stat_sq),
                       C.cast(C.pointer(bn), C.c_void_p) if bn is not None else C.c_void_p(0))
    lib = _lib.cuda_lib()
    _lib.check(lib.idcb_conv_fprop(C.byref(d), _lib.ptr(x), _lib.ptr(w), _lib.ptr(y), C.byref(f),
                                   _lib.stream()))
    _lib.count_launch()


def conv_dgrad_raw(g: ConvGeom, dy: torch.Tensor, w: torch.Tensor, dx: torch.Tensor | None, *,
                   dy_coff: int = 0, dx_coff: int = 0, dx_pitch: int | None = None, mode: int = 0,
                   aux=None, aux_coff: int = 0, bwd_scale=None, bwd_shift=None, bwd_act: int = ACT_NONE,
                   gacc=None, g_coff: int = 0, stat_a=None, stat_b=None, tail=None, slice_tail=None) -> None:
    """dx = epi(conv^T(dy, w)); see include/idcb200.h for the epilogue modes. ``tail`` (bn_tail(...)): finish
    the BatchNorm backward of the mode-2 epilogue in the same launch."""
    _chk(dy, torch.bfloat16, "dy"); _chk(w, torch.bfloat16, "w"); _chk(dx, torch.bfloat16, "dx")
    _chk(aux, torch.bfloat16, "aux"); _chk(gacc, torch.float32, "gacc")
    b = dy.shape[0]
    if dx_pitch is None:
        dx_pitch = dx.shape[-1] if dx is not None else g.cin
    d = _desc(g, b, dx_pitch, dx_coff, dy.shape[-1], dy_coff)
    f = _lib.ConvBwdFused(mode, _lib.ptr(aux), aux.shape[-1] if aux is not None else 0, aux_coff,
                          _lib.ptr(bwd_scale), _lib.ptr(bwd_shift), bwd_act, _lib.ptr(gacc),
                          gacc.shape[-1] if gacc is not None else 0, g_coff, _lib.ptr(stat_a),
                          _lib.ptr(stat_b),
                          C.cast(C.pointer(tail), C.c_void_p) if tail is not None else C.c_void_p(0),
                          C.cast(C.pointer(slice_tail), C.c_void_p) if slice_tail is not None else C.c_void_p(0))
    lib = _lib.cuda_lib()
    _lib.check(lib.idcb_conv_dgrad(C.byref(d), _lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), C.byref(f),
                                   _lib.stream()))
    _lib.count_launch()


def conv_wgrad_raw(g: ConvGeom, x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, *, x_coff: int = 0,
                   dy_coff: int = 0, pro_scale=None, pro_shift=None, pro_act: int = ACT_NONE) -> None:
    """dw (fp32 [Cout,KH,KW,Cin]) += pro(x)^T * dy."""
    _chk(x, torch.bfloat16, "x"); _chk(dy, torch.bfloat16, "dy"); _chk(dw, torch.float32, "dw")
    b = x.shape[0]
    d = _desc(g, b, x.shape[-1], x_coff, dy.shape[-1], dy_coff)
    lib = _lib.cuda_lib()
    _lib.check(lib.idcb_conv_wgrad(C.byref(d), _lib.ptr(x), _lib.ptr(dy), _lib.ptr(dw),
                                   _lib.ptr(pro_scale), _lib.ptr(pro_shift), pro_act, _lib.stream()))
    _lib.count_launch()


def bn_tail(out: torch.Tensor, mean, rstd, count, counter=None, dgamma=None, dbeta=None, out_coff: int = 0,
            frozen: bool = False) -> _lib.BnTail:
    """BatchNorm backward to be finished inside a mode-2 dgrad launch (include/idcb200.h: idcb_bn_tail).
    ``counter``: a zeroed 4-byte slot (any dtype) private to this launch."""
    p = _lib.ptr
    return _lib.BnTail(p(out), out.shape[-1], out_coff, p(mean), p(rstd), float(count), p(dgamma), p(dbeta),
                       p(counter), 1 if frozen else 0)


def slice_tail(out: torch.Tensor, c0: int, acc_s1, acc_s2, mean, rstd, bn_mean, bn_rstd, count, counter,
               dgamma=None, dbeta=None, out_coff: int = 0) -> _lib.SliceTail:
    """The finished gradient of block-buffer channels [c0, c0 + 32) produced inside the mode-2 dgrad launch that
    writes them last (include/idcb200.h: idcb_slice_tail)."""
    p = _lib.ptr
    return _lib.SliceTail(p(out), out.shape[-1], out_coff, p(acc_s1), p(acc_s2), p(mean), p(rstd), p(bn_mean),
                          p(bn_rstd), p(dgamma), p(dbeta), float(count), p(counter), int(c0))


def slice_tail_fits(rows: int, cin: int) -> bool:
    """co-residency rule of the slice tail: 128-row x 32-column tiles, two CTAs per SM"""
    sms = _lib.cuda_lib().idcb_device_sm_count()
    return -(-rows // 128) * -(-cin // 32) <= 2 * sms


def bn_inline(sum_, sq, gamma, beta, count, eps, momentum, scale_out, shift_out, mean_out, rstd_out,
              moving_mean=None, moving_var=None) -> _lib.BnInline:
    """Description of a training-mode BatchNorm to be finalised inside the consuming conv kernel."""
    p = _lib.ptr
    return _lib.BnInline(p(sum_), p(sq), p(gamma), p(beta), float(count), float(eps), float(momentum),
                         p(scale_out), p(shift_out), p(mean_out), p(rstd_out), p(moving_mean),
                         p(moving_var))


# ---------------------------------------------------------------------------------------------
# PyTorch fp32 statement of the same math (CPU execution path and numerical oracle)
# ---------------------------------------------------------------------------------------------
def act_ref(v: torch.Tensor, act: int) -> torch.Tensor:
    if act == ACT_RELU:
        return torch.relu(v)
    if act == ACT_RELU6:
        return torch.clamp(v, 0.0, 6.0)
    return v


def conv_ref(g: ConvGeom, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """NHWC conv with explicit (possibly asymmetric) zero padding. x [B,H,W,Cin], w [Cout,KH,KW,Cin]."""
    xn = x.permute(0, 3, 1, 2)
    xn = F.pad(xn, (g.pad_l, g.pad_r, g.pad_t, g.pad_b))
    y = F.conv2d(xn, w.permute(0, 3, 1, 2), stride=g.stride)
    return y.permute(0, 2, 3, 1).contiguous()
