"""Fused sm_100a pipeline for MobileNetV2 (alpha 1.0) + GAP + Dense head
(/root/reference/dist_model_tf_mobile.py:119-129; FedAvg north-star config uses it too).

Per inverted-residual block: expand 1x1 (tcgen05) -> depthwise 3x3 (direct kernel) -> project 1x1
(tcgen05) -> one element-wise kernel that applies the project BatchNorm and the residual Add. Every
other BatchNorm + ReLU6 is a per-channel scale/shift applied by the CONSUMING kernel while it loads its
operand, and each producer accumulates the batch statistics of what it stores, so the activations of
the 52 BatchNorm / 35 ReLU6 layers never make a round trip through memory. Keras ``correct_pad``
asymmetric padding of the stride-2 depthwise convs and of ``Conv1`` is expressed as (pad_t, pad_l).
"""
from __future__ import annotations

import os

import torch

from ...ops import conv as OC
from ...ops import kernels as K
from ..spec import MBV2_CFG, correct_pad
from .common import ACT_NONE, ACT_RELU6, BNState, FusedExecutor, Scratch


class _Plan:
    pass


class _Blk:
    pass


class _Stat:
    """sum / sum-of-squares slots in the per-step scratch for one produced tensor"""

    def __init__(self, sc: Scratch, c: int):
        self.sc, self.c = sc, c
        self.o_sum, self.o_sq = sc.reserve(c), sc.reserve(c)

    @property
    def sum(self):
        return self.sc.view(self.o_sum, self.c)

    @property
    def sq(self):
        return self.sc.view(self.o_sq, self.c)


class MobileNetV2Executor(FusedExecutor):
    def build_plan(self, batch: int):
        net = self.net
        dev = net.device
        bf = torch.bfloat16
        ft = self.first_trainable
        p = _Plan()
        H, W, C = net.input_shape
        p.x32 = torch.zeros(batch, H, W, C, dtype=torch.float32, device=dev)
        p.labels = torch.zeros(batch, dtype=torch.float32, device=dev)
        p.acc = torch.zeros(4, dtype=torch.float32, device=dev)
        p.logits = torch.zeros(batch, net.n_outputs, dtype=torch.float32, device=dev)
        p.x16 = torch.zeros(batch, H, W, 8, dtype=bf, device=dev)
        sc = Scratch(dev)
        p.scratch = sc

        def idx(n):
            return net.get_layer(n).index

        def z(h, w, c, dtype=bf):
            return torch.zeros(batch, h, w, c, dtype=dtype, device=dev)

        def coef(c):
            return tuple(torch.zeros(c, device=dev) for _ in range(3))

        # stem: Conv1_pad + Conv1 3x3/2 VALID
        p.g_stem = OC.geom(H, W, 8, 32, 3, 2, (correct_pad(H), correct_pad(W)))
        h, w = p.g_stem.ho, p.g_stem.wo
        p.s_raw = z(h, w, 32)
        p.st_stem = _Stat(sc, 32)
        p.i_stem, p.i_stem_bn = idx("Conv1"), idx("bn_Conv1")
        p.bn_stem = BNState(self, "bn_Conv1", 32, sc, ft <= p.i_stem_bn)
        if ft <= p.i_stem_bn:
            p.dpre_stem = z(h, w, 32)
            p.ds_raw = z(h, w, 32)
            p.c_stem = coef(32)
        p.blocks = []
        cin = 32
        block_id = 0
        from ..spec import MBV2_CFGS
        for t, c, n, s in MBV2_CFGS[net.base.name]:
            for i in range(n):
                stride = s if i == 0 else 1
                b = _Blk()
                b.prefix = "expanded_conv_" if block_id == 0 else f"block_{block_id}_"
                b.has_expand = block_id != 0
                b.stride, b.cin, b.cout, b.cexp = stride, cin, c, (t * cin if block_id != 0 else cin)
                b.add = cin == c and stride == 1
                b.h, b.w = h, w
                if b.has_expand:
                    b.i_e, b.i_ebn = idx(b.prefix + "expand"), idx(b.prefix + "expand_BN")
                    b.ge = OC.geom(h, w, cin, b.cexp, 1)
                    b.e_raw = z(h, w, b.cexp)
                    b.st_e = _Stat(sc, b.cexp)
                    b.bn_e = BNState(self, b.prefix + "expand_BN", b.cexp, sc, ft <= b.i_ebn)
                b.i_d, b.i_dbn = idx(b.prefix + "depthwise"), idx(b.prefix + "depthwise_BN")
                pad = "same" if stride == 1 else (correct_pad(h), correct_pad(w))
                b.gd = OC.geom(h, w, b.cexp, b.cexp, 3, stride, pad)
                ho, wo = b.gd.ho, b.gd.wo
                b.d_raw = z(ho, wo, b.cexp)
                b.st_d = _Stat(sc, b.cexp)
                b.bn_d = BNState(self, b.prefix + "depthwise_BN", b.cexp, sc, ft <= b.i_dbn)
                b.o_ctr = sc.reserve(64)        # grid-barrier word of the project dgrad + BatchNorm-tail launch
                b.i_p, b.i_pbn = idx(b.prefix + "project"), idx(b.prefix + "project_BN")
                b.gp = OC.geom(ho, wo, b.cexp, c, 1)
                b.p_raw = z(ho, wo, c)
                b.st_p = _Stat(sc, c)
                b.bn_p = BNState(self, b.prefix + "project_BN", c, sc, ft <= b.i_pbn)
                b.out = z(ho, wo, c)
                b.ho, b.wo = ho, wo
                # backward buffers
                b.bwd = ft <= b.i_pbn
                if b.bwd:
                    b.dout = z(ho, wo, c)            # dL/d(out), written by the consumer
                    b.dp_raw = z(ho, wo, c)
                    b.c_p = coef(c)
                    if ft <= b.i_dbn:
                        b.dpre_d = z(ho, wo, b.cexp)
                        b.dd_raw = z(ho, wo, b.cexp)
                        b.c_d = coef(b.cexp)
                    if b.has_expand and ft <= b.i_ebn:
                        b.dpre_e = z(h, w, b.cexp)
                        b.de_raw = z(h, w, b.cexp)
                        b.c_e = coef(b.cexp)
                p.blocks.append(b)
                cin, h, w = c, ho, wo
                block_id += 1
        # Conv_1 1x1 -> 1280
        p.i_c, p.i_cbn = idx("Conv_1"), idx("Conv_1_bn")
        p.gc = OC.geom(h, w, cin, 1280, 1)
        p.c_raw = z(h, w, 1280)
        p.st_c = _Stat(sc, 1280)
        p.bn_c = BNState(self, "Conv_1_bn", 1280, sc, ft <= p.i_cbn)
        p.hw = h * w
        if ft <= p.i_cbn:
            p.dpre_c = z(h, w, 1280)
            p.dc_raw = z(h, w, 1280)
            p.c_c = coef(1280)
        p.head_layer = net.layers[-1]
        sc.finalize()
        return p

    # -------------------------------------------------------------------------------------------------
    def _bn(self, bn: BNState, training: bool, st: _Stat, count: float) -> None:
        if training and bn.trainable:
            bn.finalize(st.sum, st.sq, count)
        else:
            bn.fold()

    def _pro(self, bn: BNState, training: bool, st: _Stat, count: float) -> dict:
        """kwargs that make ``bn`` + activation the operand prologue of the consuming tensor-core conv: a training-mode
        BatchNorm is finalised INSIDE that kernel from the batch sums (no finalisation launch)."""
        if training and bn.trainable:
            bn._folded_at = None
            return {"bn": OC.bn_inline(st.sum, st.sq, bn.gamma, bn.beta, count, bn.eps, bn.momentum, bn.scale,
                                       bn.shift, bn.mean, bn.rstd, bn.mmean, bn.mvar)}
        bn.fold()
        return {"pro_scale": bn.scale, "pro_shift": bn.shift}

    @staticmethod
    def _cnt(t: torch.Tensor) -> float:
        return float(t.numel() // t.shape[-1])

    def run(self, p, training: bool, grad_scale: float) -> None:
        net = self.net
        p.scratch.zero()
        p.acc.zero_()
        tr = training

        def ss(st):
            return (st.sum, st.sq) if tr else (None, None)

        K.pack_input(p.x32, p.x16)
        a, q = ss(p.st_stem)
        OC.conv_fprop_raw(p.g_stem, p.x16, self.w16("Conv1/kernel"), p.s_raw, stat_sum=a, stat_sq=q)
        self._bn(p.bn_stem, tr, p.st_stem, self._cnt(p.s_raw))
        cur = None      # materialised block input; None -> lazy stem output
        for b in p.blocks:
            if b.has_expand:
                a, q = ss(b.st_e)
                OC.conv_fprop_raw(b.ge, cur, self.w16(b.prefix + "expand/kernel"), b.e_raw, stat_sum=a, stat_sq=q)
                self._bn(b.bn_e, tr, b.st_e, self._cnt(b.e_raw))
                dw_in, dw_bn = b.e_raw, b.bn_e
            else:
                dw_in, dw_bn = p.s_raw, p.bn_stem
            a, q = ss(b.st_d)
            K.dwconv_fprop(b.gd, dw_in, self.w16(b.prefix + "depthwise/depthwise_kernel"), b.d_raw,
                           scale=dw_bn.scale, shift=dw_bn.shift, act=ACT_RELU6, stat_sum=a, stat_sq=q)
            a, q = ss(b.st_p)
            OC.conv_fprop_raw(b.gp, b.d_raw, self.w16(b.prefix + "project/kernel"), b.p_raw, pro_act=ACT_RELU6,
                              stat_sum=a, stat_sq=q, **self._pro(b.bn_d, tr, b.st_d, self._cnt(b.d_raw)))
            self._bn(b.bn_p, tr, b.st_p, self._cnt(b.p_raw))
            K.affine_act(b.p_raw, b.out, b.cout, b.bn_p.scale, b.bn_p.shift, ACT_NONE,
                         res=cur if b.add else None)
            cur = b.out
        a, q = ss(p.st_c)
        OC.conv_fprop_raw(p.gc, cur, self.w16("Conv_1/kernel"), p.c_raw, stat_sum=a, stat_sq=q)
        self._bn(p.bn_c, tr, p.st_c, self._cnt(p.c_raw))
        hl = p.head_layer
        wname, bname = hl.spec.params[0].name, hl.spec.params[1].name
        head_train = tr and hl.index in self.trainable
        bwd = tr and self.first_trainable <= p.i_cbn
        K.head(p.c_raw, p.hw, 1280, net.n_outputs, self.w32(wname), self.w32(bname), p.labels, p.logits,
               p.acc, pro_scale=p.bn_c.scale, pro_shift=p.bn_c.shift, pro_act=ACT_RELU6,
               loss_kind=self.loss_kinds[net.loss.kind], acc_on_prob=net.acc_on_prob,
               threshold=net.acc_threshold, training=tr, grad_scale=grad_scale,
               dw=self.gw(wname) if head_train else None, db=self.gw(bname) if head_train else None,
               dx=p.dpre_c if bwd else None, stat_a=p.bn_c.sa if bwd else None,
               stat_b=p.bn_c.sb if bwd else None)
        if bwd:
            self._backward(p)

    # -------------------------------------------------------------------------------------------------
    def _bn_bwd(self, bn: BNState, count: float, coefs, need_dx: bool) -> None:
        """sa/sb of ``bn`` are filled. If the input gradient is needed, the parameter gradients are
        emitted by the fused apply kernel that follows (:meth:`_bn_apply`); otherwise only dgamma/dbeta."""
        if need_dx:
            self._pending = (bn, count)
        elif bn.trainable:
            K.bn_bwd_coeffs(bn.sa, bn.sb, bn.mean, bn.rstd, bn.scale, count, dgamma=bn.dgamma,
                            dbeta=bn.dbeta)

    def _bn_apply(self, g, x, out, channels) -> None:
        bn, count = self._pending
        K.bn_bwd_apply(0 if bn.trainable else 1, g, x, out, channels, count, sa=bn.sa, sb=bn.sb,
                       mean=bn.mean, rstd=bn.rstd, scale=bn.scale, dgamma=bn.dgamma, dbeta=bn.dbeta)

    # Weight-gradient kernels (1x1 and depthwise) run on a side stream beside the data-gradient chain: every
    # buffer they read is per-block and written once per step, so the only join is at the end of the backward.
    wgrad_side_stream = True
    bn_tail = os.environ.get("IDCB_BN_TAIL", "1") != "0"

    def _side(self, fn, *args, **kw) -> None:
        if not self.wgrad_side_stream:
            fn(*args, **kw)
            return
        if getattr(self, "_wg_stream", None) is None:
            self._wg_stream = torch.cuda.Stream(device=self.net.device)
        self._wg_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._wg_stream):
            fn(*args, **kw)
        self._wg_pending = True

    # ---- gradient announcements: once the backward pass has left inverted-residual block bi, every
    # gradient of block bi and of everything above it is final, so that part of the flat gradient buffer can be
    # all-reduced while the blocks below are still being differentiated
    def _block_offsets_table(self) -> list:
        st = self.net.store
        offs = []
        block_id = 0
        while True:
            prefix = "expanded_conv_" if block_id == 0 else f"block_{block_id}_"
            names = [pp.name for l in self.net.layers if l.spec.name.startswith(prefix) for pp in l.spec.params
                     if st.slots[pp.name].trainable_kind]
            if not names:
                break
            offs.append(min(st.slots[n].offset for n in names))
            block_id += 1
        return offs

    def grad_ready_offsets(self) -> list:
        self._blk_off = self._block_offsets_table()
        return sorted(self._blk_off, reverse=True)

    def _announce(self, bi: int) -> None:
        if not hasattr(self, "_blk_off"):
            self._blk_off = self._block_offsets_table()
        if bi >= len(self._blk_off):
            return
        off = self._blk_off[bi]
        if not self.announcement_fires(off):
            return          # merged into a later bucket: no need to wait for the weight-gradient stream now
        if getattr(self, "_wg_pending", False):
            torch.cuda.current_stream().wait_stream(self._wg_stream)
            self._wg_pending = False
        self.grads_ready(off)

    def _backward(self, p) -> None:
        try:
            self._backward_impl(p)
        finally:
            if getattr(self, "_wg_pending", False):
                torch.cuda.current_stream().wait_stream(self._wg_stream)
                self._wg_pending = False

    def _backward_impl(self, p) -> None:
        ft, tr = self.first_trainable, self.trainable
        # ---- Conv_1 -------------------------------------------------------------------------------------
        last = p.blocks[-1]
        need = ft <= p.i_c
        self._bn_bwd(p.bn_c, self._cnt(p.c_raw), p.c_c, need)
        if not need:
            return
        self._bn_apply(p.dpre_c, p.c_raw, p.dc_raw, 1280)
        if p.i_c in tr:
            self._side(OC.conv_wgrad_raw, p.gc, last.out, p.dc_raw, self.gw("Conv_1/kernel"))
        if not last.bwd:
            return
        OC.conv_dgrad_raw(p.gc, p.dc_raw, self.w16("Conv_1/kernel"), last.dout, mode=0)
        # ---- blocks -------------------------------------------------------------------------------------
        for bi in range(len(p.blocks) - 1, -1, -1):
            b = p.blocks[bi]
            if not b.bwd:
                return
            prev = p.blocks[bi - 1] if bi > 0 else None
            xin = prev.out if prev is not None else None
            # project BN (no activation): dpre = dout
            K.channel_stats2(b.dout, b.p_raw, b.cout, b.bn_p.sa, b.bn_p.sb)
            need = ft <= b.i_p
            self._bn_bwd(b.bn_p, self._cnt(b.p_raw), b.c_p, need)
            if not need:
                # a residual branch may still carry gradient further down
                if b.add and prev is not None and prev.bwd:
                    prev.dout.copy_(b.dout)
                    continue
                return
            self._bn_apply(b.dout, b.p_raw, b.dp_raw, b.cout)
            if b.i_p in tr:
                self._side(OC.conv_wgrad_raw, b.gp, b.d_raw, b.dp_raw, self.gw(b.prefix + "project/kernel"),
                                  pro_scale=b.bn_d.scale, pro_shift=b.bn_d.shift, pro_act=ACT_RELU6)
            stop = ft > b.i_dbn
            if not stop:
                need = ft <= b.i_d
                # the depthwise BatchNorm's backward rides in the project dgrad launch (column sums, grid barrier, the
                # tile parked in TMEM); the native side falls back to dgrad + bn_bwd_apply where the grid has more tiles
                # than SMs
                tail = None
                if need and self.bn_tail:
                    tail = OC.bn_tail(b.dd_raw, b.bn_d.mean, b.bn_d.rstd, self._cnt(b.d_raw),
                                      counter=p.scratch.view(b.o_ctr, 1), dgamma=b.bn_d.dgamma, dbeta=b.bn_d.dbeta,
                                      frozen=not b.bn_d.trainable)
                OC.conv_dgrad_raw(b.gp, b.dp_raw, self.w16(b.prefix + "project/kernel"), b.dpre_d, mode=2,
                                  aux=b.d_raw, bwd_scale=b.bn_d.scale, bwd_shift=b.bn_d.shift,
                                  bwd_act=ACT_RELU6, stat_a=b.bn_d.sa, stat_b=b.bn_d.sb, tail=tail)
                if tail is None:
                    self._bn_bwd(b.bn_d, self._cnt(b.d_raw), b.c_d, need)
                stop = not need
            if not stop:
                if tail is None:
                    self._bn_apply(b.dpre_d, b.d_raw, b.dd_raw, b.cexp)
                dw_in, dw_bn = (b.e_raw, b.bn_e) if b.has_expand else (p.s_raw, p.bn_stem)
                if b.i_d in tr:
                    self._side(K.dwconv_wgrad, b.gd, dw_in, b.dd_raw, self.gw(b.prefix + "depthwise/depthwise_kernel"),
                                   scale=dw_bn.scale, shift=dw_bn.shift, act=ACT_RELU6)
                below = ft <= (b.i_ebn if b.has_expand else p.i_stem_bn)
                if below:
                    dpre = b.dpre_e if b.has_expand else p.dpre_stem
                    K.dwconv_dgrad(b.gd, b.dd_raw, self.w16(b.prefix + "depthwise/depthwise_kernel"), dw_in,
                                   dpre, scale=dw_bn.scale, shift=dw_bn.shift, act=ACT_RELU6,
                                   stat_a=dw_bn.sa, stat_b=dw_bn.sb)
                    if b.has_expand:
                        need = ft <= b.i_e
                        self._bn_bwd(b.bn_e, self._cnt(b.e_raw), b.c_e, need)
                        if need:
                            self._bn_apply(b.dpre_e, b.e_raw, b.de_raw, b.cexp)
                            if b.i_e in tr:
                                self._side(OC.conv_wgrad_raw, b.ge, xin, b.de_raw, self.gw(b.prefix + "expand/kernel"))
                            if prev is not None and prev.bwd:
                                OC.conv_dgrad_raw(b.ge, b.de_raw, self.w16(b.prefix + "expand/kernel"),
                                                  prev.dout, mode=0, aux=b.dout if b.add else None)
                                self._announce(bi)
                                continue
                    else:
                        # first block: the depthwise conv reads the stem directly
                        need = ft <= p.i_stem
                        self._bn_bwd(p.bn_stem, self._cnt(p.s_raw), p.c_stem, need)
                        if need:
                            self._bn_apply(p.dpre_stem, p.s_raw, p.ds_raw, 32)
                            self._side(OC.conv_wgrad_raw, p.g_stem, p.x16, p.ds_raw, self.gw("Conv1/kernel"))
                        return
            # the main branch stopped inside this block; only a residual can carry gradient on
            if b.add and prev is not None and prev.bwd:
                prev.dout.copy_(b.dout)
                continue
            return
