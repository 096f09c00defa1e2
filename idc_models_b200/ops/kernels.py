"""Thin Python bindings for the non-GEMM kernels of ``libidcb200.so`` (csrc/elementwise.cu,
depthwise.cu, head.cu). Tensors are NHWC bf16 activations ``[B, H, W, pitch]`` or flat fp32 vectors;
every call is asynchronous on the current CUDA stream. See include/idcb200.h for the contracts and
the reference citations.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .conv import ConvGeom, _desc

vp = C.c_void_p


class PoolDesc(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("batch", "h", "w", "c", "ho", "wo", "k", "stride", "pad", "is_max",
                                       "x_pitch", "x_coff", "y_pitch", "y_coff")]


class HeadDesc(C.Structure):
    _fields_ = [("batch", C.c_int), ("hw", C.c_int), ("c", C.c_int), ("k", C.c_int),
                ("x_pitch", C.c_int), ("x_coff", C.c_int),
                ("pro_scale", vp), ("pro_shift", vp), ("pro_act", C.c_int),
                ("loss_kind", C.c_int), ("acc_on_prob", C.c_int), ("threshold", C.c_float),
                ("training", C.c_int), ("grad_scale", C.c_float),
                ("dx_pitch", C.c_int), ("dx_coff", C.c_int), ("g_pitch", C.c_int), ("g_coff", C.c_int)]


def _p(t):
    return _lib.ptr(t)


def _call(fn, *args):
    _lib.check(fn(*args, _lib.stream()))
    _lib.count_launch()


def _off(t: torch.Tensor | None, off: int):
    """pointer to element `off` of a flat fp32 tensor (None -> NULL)"""
    if t is None:
        return vp(0)
    return vp(t.data_ptr() + off * t.element_size())


def bn_finalize(sum_, sq, count, gamma, beta, eps, scale, shift, mean_out=None, rstd_out=None,
                moving_mean=None, moving_var=None, momentum=0.99):
    c = scale.numel()
    _call(_lib.cuda_lib().idcb_bn_finalize, _p(sum_), _p(sq), C.c_float(count), _p(gamma), _p(beta),
          C.c_float(eps), _p(scale), _p(shift), _p(mean_out), _p(rstd_out), _p(moving_mean),
          _p(moving_var), C.c_float(momentum), C.c_int(c))


def bn_fold(gamma, beta, mean, var, eps, scale, shift):
    _call(_lib.cuda_lib().idcb_bn_fold, _p(gamma), _p(beta), _p(mean), _p(var), C.c_float(eps),
          _p(scale), _p(shift), C.c_int(scale.numel()))


def bn_bwd_coeffs(sa, sb, mean, rstd, scale, count, dgamma=None, dbeta=None, A=None, B=None, C0=None,
                  acc_s1=None, acc_s2=None):
    _call(_lib.cuda_lib().idcb_bn_bwd_coeffs, _p(sa), _p(sb), _p(mean), _p(rstd), _p(scale),
          C.c_float(count), _p(dgamma), _p(dbeta), _p(A), _p(B), _p(C0), _p(acc_s1), _p(acc_s2),
          C.c_int(sa.numel()))


def bn_shared_coeffs(s1, s2, mean, rstd, count, A, B, C0):
    _call(_lib.cuda_lib().idcb_bn_shared_coeffs, _p(s1), _p(s2), _p(mean), _p(rstd), C.c_float(count),
          _p(A), _p(B), _p(C0), C.c_int(s1.numel()))


def _rows(t: torch.Tensor) -> int:
    return t.numel() // t.shape[-1]


def affine_act(x, y, channels, scale=None, shift=None, act=0, x_coff=0, y_coff=0, res=None, res_coff=0):
    _call(_lib.cuda_lib().idcb_affine_act, _p(x), C.c_int(x.shape[-1]), C.c_int(x_coff), _p(y),
          C.c_int(y.shape[-1]), C.c_int(y_coff), _p(scale), _p(shift), C.c_int(act), _p(res),
          C.c_int(res.shape[-1] if res is not None else 8), C.c_int(res_coff), C.c_longlong(_rows(x)),
          C.c_int(channels))


def affine2(g, x, out, channels, A, B, C0, g_coff=0, x_coff=0, o_coff=0):
    _call(_lib.cuda_lib().idcb_affine2, _p(g), C.c_int(1 if g.dtype == torch.float32 else 0),
          C.c_int(g.shape[-1]), C.c_int(g_coff), _p(x), C.c_int(x.shape[-1]), C.c_int(x_coff), _p(out),
          C.c_int(out.shape[-1]), C.c_int(o_coff), _p(A), _p(B), _p(C0), C.c_longlong(_rows(x)),
          C.c_int(channels))


def bn_bwd_apply(mode, g, x, out, channels, count, *, sa=None, sb=None, mean=None, rstd=None, scale=None,
                 dgamma=None, dbeta=None, g_coff=0, x_coff=0, o_coff=0):
    """BatchNorm backward in one launch (see include/idcb200.h): mode 0 training BN, 1 frozen BN,
    2 DenseNet shared-buffer slice (sa=S1, sb=S2)."""
    _call(_lib.cuda_lib().idcb_bn_bwd_apply, C.c_int(mode), _p(g),
          C.c_int(1 if g.dtype == torch.float32 else 0), C.c_int(g.shape[-1]), C.c_int(g_coff), _p(x),
          C.c_int(x.shape[-1]), C.c_int(x_coff), _p(out), C.c_int(out.shape[-1]), C.c_int(o_coff), _p(sa),
          _p(sb), _p(mean), _p(rstd), _p(scale), C.c_float(count), _p(dgamma), _p(dbeta),
          C.c_longlong(_rows(x)), C.c_int(channels))


def bn_bwd_slice(g, x, out, c_lo, channels, count, acc_s1, acc_s2, mean, rstd, pending=None):
    """DenseNet block-buffer slice backward with the pending consumer BatchNorm's bookkeeping folded in.
    ``pending`` = (sa, sb, mean, rstd, scale, dgamma, dbeta) over channels [0, len(sa)) of the block."""
    ps = pending or (None,) * 7
    _call(_lib.cuda_lib().idcb_bn_bwd_slice, _p(g), C.c_int(g.shape[-1]), C.c_int(c_lo), _p(x),
          C.c_int(x.shape[-1]), C.c_int(c_lo), _p(out), C.c_int(out.shape[-1]), C.c_int(0), _p(acc_s1), _p(acc_s2),
          C.c_int(c_lo), _off(mean, c_lo), _off(rstd, c_lo), C.c_float(count), C.c_longlong(_rows(x)),
          C.c_int(channels), _p(ps[0]), _p(ps[1]), _p(ps[2]), _p(ps[3]), _p(ps[4]), _p(ps[5]), _p(ps[6]),
          C.c_int(ps[0].numel() if ps[0] is not None else 0))


def channel_stats(x, channels, sum_, sq, coff=0):
    _call(_lib.cuda_lib().idcb_channel_stats, _p(x), C.c_int(x.shape[-1]), C.c_int(coff),
          C.c_longlong(_rows(x)), C.c_int(channels), _p(sum_), _p(sq))


def channel_stats2(g, x, channels, sum_g, sum_gx, g_coff=0, x_coff=0):
    _call(_lib.cuda_lib().idcb_channel_stats2, _p(g), C.c_int(g.shape[-1]), C.c_int(g_coff), _p(x),
          C.c_int(x.shape[-1]), C.c_int(x_coff), C.c_longlong(_rows(x)), C.c_int(channels), _p(sum_g),
          _p(sum_gx))


def pool_desc(x, y, channels, k, stride, pad, is_max, x_coff=0, y_coff=0) -> PoolDesc:
    b, h, w, xp = x.shape
    _, ho, wo, yp = y.shape
    return PoolDesc(b, h, w, channels, ho, wo, k, stride, pad, 1 if is_max else 0, xp, x_coff, yp, y_coff)


def pool_fwd(x, y, channels, k, stride, pad=0, is_max=True, x_coff=0, y_coff=0, scale=None, shift=None, act=0,
             argmax=None):
    """``argmax`` (uint8 [B, Ho, Wo, channels], max pooling): records the winning window element for pool_bwd"""
    d = pool_desc(x, y, channels, k, stride, pad, is_max, x_coff, y_coff)
    _call(_lib.cuda_lib().idcb_pool_fwd, C.byref(d), _p(x), _p(y), _p(scale), _p(shift), C.c_int(act), _p(argmax))


def pool_bwd(x, y, dy, dx, channels, k, stride, pad=0, is_max=True, x_coff=0, y_coff=0, dy_coff=0,
             dx_coff=0, scale=None, shift=None, act=0, relu_mask=False, stat_a=None, stat_b=None, argmax=None):
    d = pool_desc(x, y, channels, k, stride, pad, is_max, x_coff, y_coff)
    _call(_lib.cuda_lib().idcb_pool_bwd, C.byref(d), _p(x), _p(y), _p(dy), C.c_int(dy.shape[-1]),
          C.c_int(dy_coff), _p(dx), C.c_int(dx.shape[-1]), C.c_int(dx_coff), _p(scale), _p(shift),
          C.c_int(act), C.c_int(1 if relu_mask else 0), _p(stat_a), _p(stat_b), _p(argmax))


def pack_input(x_f32: torch.Tensor, y_bf16: torch.Tensor):
    """fp32 NHWC [B,H,W,Cin] -> bf16 NHWC [B,H,W,Cpad] with zero pad channels."""
    cin, cpad = x_f32.shape[-1], y_bf16.shape[-1]
    _call(_lib.cuda_lib().idcb_pack_input, _p(x_f32), _p(y_bf16), C.c_longlong(_rows(x_f32)),
          C.c_int(cin), C.c_int(cpad))


def cast_bf16(x_f32, y_bf16):
    _call(_lib.cuda_lib().idcb_cast_f32_bf16, _p(x_f32), _p(y_bf16), C.c_longlong(x_f32.numel()))


def head(x, hw, channels, k, w, bias, labels, logits, acc, *, x_coff=0, pro_scale=None, pro_shift=None,
         pro_act=0, loss_kind=0, acc_on_prob=False, threshold=0.5, training=False, grad_scale=1.0,
         dw=None, db=None, dx=None, dx_coff=0, gacc=None, g_coff=0, stat_a=None, stat_b=None):
    batch = labels.numel()
    d = HeadDesc(batch, hw, channels, k, x.shape[-1], x_coff, _p(pro_scale), _p(pro_shift), pro_act,
                 loss_kind, 1 if acc_on_prob else 0, threshold, 1 if training else 0, grad_scale,
                 dx.shape[-1] if dx is not None else 0, dx_coff,
                 gacc.shape[-1] if gacc is not None else 0, g_coff)
    _call(_lib.cuda_lib().idcb_head, C.byref(d), _p(x), _p(w), _p(bias), _p(labels), _p(logits), _p(acc),
          _p(dw), _p(db), _p(dx), _p(gacc), _p(stat_a), _p(stat_b))


def dwconv_fprop(g: ConvGeom, x, w, y, *, x_coff=0, y_coff=0, scale=None, shift=None, act=0,
                 stat_sum=None, stat_sq=None):
    d = _desc(g, x.shape[0], x.shape[-1], x_coff, y.shape[-1], y_coff)
    _call(_lib.cuda_lib().idcb_dwconv_fprop, C.byref(d), _p(x), _p(w), _p(y), _p(scale), _p(shift),
          C.c_int(act), _p(stat_sum), _p(stat_sq))


def dwconv_dgrad(g: ConvGeom, dy, w, x, dx, *, scale=None, shift=None, act=0, stat_a=None, stat_b=None):
    d = _desc(g, dy.shape[0], dx.shape[-1], 0, dy.shape[-1], 0)
    _call(_lib.cuda_lib().idcb_dwconv_dgrad, C.byref(d), _p(dy), _p(w), _p(x), _p(dx), _p(scale),
          _p(shift), C.c_int(act), _p(stat_a), _p(stat_b))


def dwconv_wgrad(g: ConvGeom, x, dy, dw, *, scale=None, shift=None, act=0):
    d = _desc(g, x.shape[0], x.shape[-1], 0, dy.shape[-1], 0)
    _call(_lib.cuda_lib().idcb_dwconv_wgrad, C.byref(d), _p(x), _p(dy), _p(dw), _p(scale), _p(shift),
          C.c_int(act))
