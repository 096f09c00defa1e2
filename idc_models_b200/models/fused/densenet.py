"""Fused sm_100a pipeline for DenseNet-121/169/201 + GAP + Dense head
(/root/reference/dist_model_tf_dense.py:131-141; north-star config: DenseNet-121 on 50x50x3 IDC).

Memory-efficient dense blocks:

* every dense block owns ONE NHWC buffer ``[B, h, w, C_final]``; each 3x3 conv's epilogue stores its 32
  new channels at its channel offset, so Concatenate costs nothing;
* per-channel batch sums of the buffer are accumulated by the producing kernel's epilogue, so every
  BatchNorm that reads a prefix of the buffer is only a per-channel scale/shift that the consuming
  1x1 conv applies while loading its operand (BN + ReLU never touch memory);
* backward mirrors it: each consumer's dgrad epilogue applies the ReLU/BN mask and accumulates
  ``scale * dpre`` into ONE fp32 gradient buffer per block; when backward reaches the layer that
  produced a channel slice, the batch-statistics correction of all its consumers is applied at once
  (``dx = G - S1/N - xhat * S2/N``).

Per dense layer: forward 2 tensor-core kernels (+2 per-channel finalizers); backward 2 wgrad + 2 dgrad
+ 2 element-wise + 3 per-channel kernels. TensorFlow runs ~7 cuDNN/Eigen kernels forward and ~14
backward for the same layer.
"""
from __future__ import annotations

import os

import torch

from ...ops import conv as OC
from ...ops import kernels as K
from .common import ACT_NONE, ACT_RELU, BNState, FusedExecutor, Scratch

GROWTH = 32


class _Plan:
    pass


class _DenseLayer:
    pass


class _Block:
    pass


class DenseNetExecutor(FusedExecutor):
    def build_plan(self, batch: int):
        net = self.net
        dev = net.device
        bf = torch.bfloat16
        p = _Plan()
        p.batch = batch
        H, W, C = net.input_shape
        p.x32 = torch.zeros(batch, H, W, C, dtype=torch.float32, device=dev)
        p.labels = torch.zeros(batch, dtype=torch.float32, device=dev)
        p.acc = torch.zeros(4, dtype=torch.float32, device=dev)
        p.logits = torch.zeros(batch, net.n_outputs, dtype=torch.float32, device=dev)
        p.x16 = torch.zeros(batch, H, W, 8, dtype=bf, device=dev)
        sc = Scratch(dev)
        p.scratch = sc
        L = net.get_layer
        ft = self.first_trainable

        def idx(name):
            return L(name).index

        # ---- stem -----------------------------------------------------------------------------------
        p.g_stem = OC.geom(H, W, 8, 64, 7, 2, ((3, 3), (3, 3)))
        h1, w1 = p.g_stem.ho, p.g_stem.wo
        p.s0 = torch.zeros(batch, h1, w1, 64, dtype=bf, device=dev)
        p.o_stem_sum, p.o_stem_sq = sc.reserve(64), sc.reserve(64)
        p.stem_bwd = ft <= idx("conv1/bn")
        p.bn_stem = BNState(self, "conv1/bn", 64, sc, p.stem_bwd)
        h2, w2 = (h1 + 2 - 3) // 2 + 1, (w1 + 2 - 3) // 2 + 1
        p.pool1_arg = torch.zeros(batch, h2, w2, 64, dtype=torch.uint8, device=dev) if p.stem_bwd else None
        if p.stem_bwd:
            p.dpool1 = torch.zeros(batch, h2, w2, 64, dtype=bf, device=dev)
            p.dpre_stem = torch.zeros(batch, h1, w1, 64, dtype=bf, device=dev)
            p.ds0 = torch.zeros(batch, h1, w1, 64, dtype=bf, device=dev)
            p.cA = torch.zeros(64, device=dev); p.cB = torch.zeros(64, device=dev); p.cC = torch.zeros(64, device=dev)

        # ---- blocks -----------------------------------------------------------------------------------
        from ..spec import DENSENET_BLOCKS
        arch = net.base.name
        cfg = DENSENET_BLOCKS[arch]
        p.blocks = []
        h, w, c0 = h2, w2, 64
        for bi, nl in enumerate(cfg):
            blk = _Block()
            blk.h, blk.w, blk.c0, blk.nl = h, w, c0, nl
            blk.ctot = c0 + GROWTH * nl
            blk.count = float(batch * h * w)
            blk.buf = torch.zeros(batch, h, w, blk.ctot, dtype=bf, device=dev)
            blk.o_sum, blk.o_sq = sc.reserve(blk.ctot), sc.reserve(blk.ctot)
            stage = f"conv{bi + 2}"
            blk.layers = []
            for li in range(nl):
                n = f"{stage}_block{li + 1}"
                dl = _DenseLayer()
                dl.name = n
                dl.cin = c0 + GROWTH * li
                dl.i_bn0, dl.i_c1, dl.i_bn1, dl.i_c2 = idx(n + "_0_bn"), idx(n + "_1_conv"), idx(n + "_1_bn"), idx(n + "_2_conv")
                dl.g1 = OC.geom(h, w, dl.cin, 4 * GROWTH, 1)
                dl.g2 = OC.geom(h, w, 4 * GROWTH, GROWTH, 3)
                if h == 1 and w == 1:
                    dl.g2 = OC.centre_tap(dl.g2)      # a 3x3 filter over one pixel: only its centre tap is ever used
                dl.y1 = torch.zeros(batch, h, w, 4 * GROWTH, dtype=bf, device=dev)
                dl.o_sum1, dl.o_sq1 = sc.reserve(4 * GROWTH), sc.reserve(4 * GROWTH)
                dl.bwd = ft <= dl.i_c2
                dl.o_ctr = sc.reserve(64) if dl.bwd else None     # grid-barrier word of the dgrad + BN-tail launch
                dl.o_ctr2 = sc.reserve(64) if dl.bwd else None    # ... and of the 1x1 dgrad + slice-tail launch
                dl.li = li
                dl.bn0 = BNState(self, n + "_0_bn", dl.cin, sc, dl.bwd and ft <= dl.i_bn0)
                dl.bn1 = BNState(self, n + "_1_bn", 4 * GROWTH, sc, dl.bwd and ft <= dl.i_bn1)
                blk.layers.append(dl)
            last = bi == len(cfg) - 1
            tname = "bn" if last else f"pool{bi + 2}_bn"
            blk.i_tbn = idx(tname)
            blk.i_end = blk.i_tbn if last else idx(f"pool{bi + 2}_conv")
            blk.bwd = ft <= blk.i_end                       # anything in or below this block trainable
            blk.needG = ft <= blk.i_tbn                     # a gradient w.r.t. the block buffer is needed
            blk.bn_t = BNState(self, tname, blk.ctot, sc, blk.needG)
            if blk.needG:
                blk.G = torch.zeros(batch, h, w, blk.ctot, dtype=torch.float32, device=dev)
                blk.o_s1, blk.o_s2 = sc.reserve(blk.ctot), sc.reserve(blk.ctot)
                # two copies: a layer's slice gradient may be written (by the layer above's 1x1 dgrad) while the
                # previous layer's weight-gradient kernel is still reading its own on the side stream
                blk.dy3 = [torch.zeros(batch, h, w, GROWTH, dtype=bf, device=dev) for _ in range(2)]
                blk.dpre1 = torch.zeros(batch, h, w, 4 * GROWTH, dtype=bf, device=dev)
                blk.dy1 = torch.zeros(batch, h, w, 4 * GROWTH, dtype=bf, device=dev)
                blk.cA = torch.zeros(blk.ctot, device=dev)
                blk.cB = torch.zeros(blk.ctot, device=dev)
                blk.cC = torch.zeros(blk.ctot, device=dev)
                blk.c1A = torch.zeros(4 * GROWTH, device=dev)
                blk.c1B = torch.zeros(4 * GROWTH, device=dev)
                blk.c1C = torch.zeros(4 * GROWTH, device=dev)
            if not last:
                t = f"pool{bi + 2}"
                blk.i_tconv = idx(t + "_conv")
                blk.gt = OC.geom(h, w, blk.ctot, blk.ctot // 2, 1)
                blk.t = torch.zeros(batch, h, w, blk.ctot // 2, dtype=bf, device=dev)
                blk.tconv = t + "_conv/kernel"
                if ft <= blk.i_tconv:
                    blk.dt = torch.zeros(batch, h, w, blk.ctot // 2, dtype=bf, device=dev)
                    blk.dpool = torch.zeros(batch, h // 2, w // 2, blk.ctot // 2, dtype=bf, device=dev)
                h, w, c0 = h // 2, w // 2, blk.ctot // 2
            p.blocks.append(blk)
        p.head_layer = net.layers[-1]
        sc.finalize()
        return p

    # -------------------------------------------------------------------------------------------------
    def _bn_forward(self, bn: BNState, training: bool, sum_, sq, count) -> None:
        if training and bn.trainable:
            bn.finalize(sum_, sq, count)
        else:
            bn.fold()

    def _bn_prologue(self, bn: BNState, training: bool, sum_, sq, count) -> dict:
        """kwargs for conv_fprop_raw that make ``bn`` (+ activation) the conv's operand prologue.
        Training-mode BatchNorm is finalised INSIDE the conv kernel from the batch sums (no extra
        launch); frozen / inference BatchNorm is folded from the moving statistics."""
        if training and bn.trainable:
            bn._folded_at = None
            return {"bn": OC.bn_inline(sum_, sq, bn.gamma, bn.beta, count, bn.eps, bn.momentum, bn.scale,
                                       bn.shift, bn.mean, bn.rstd, bn.mmean, bn.mvar)}
        bn.fold()
        return {"pro_scale": bn.scale, "pro_shift": bn.shift}

    def run(self, p, training: bool, grad_scale: float) -> None:
        net = self.net
        sc = p.scratch
        sc.zero()
        p.acc.zero_()
        stats = training  # batch statistics are only needed while training
        K.pack_input(p.x32, p.x16)
        # ---- stem: 7x7/2 conv (tensor cores, RGB padded to 8 channels) + BN + ReLU + pad + 3x3/2 max-pool
        OC.conv_fprop_raw(p.g_stem, p.x16, self.w16("conv1/conv/kernel"), p.s0,
                          stat_sum=sc.view(p.o_stem_sum, 64) if stats else None,
                          stat_sq=sc.view(p.o_stem_sq, 64) if stats else None)
        self._bn_forward(p.bn_stem, training, sc.view(p.o_stem_sum, 64), sc.view(p.o_stem_sq, 64),
                         float(p.s0.numel() // 64))
        b0 = p.blocks[0]
        K.pool_fwd(p.s0, b0.buf, 64, 3, 2, 1, True, scale=p.bn_stem.scale, shift=p.bn_stem.shift, act=ACT_RELU,
                   argmax=p.pool1_arg if training else None)
        if stats:
            K.channel_stats(b0.buf, 64, sc.view(b0.o_sum, 64), sc.view(b0.o_sq, 64))
        # ---- dense blocks ----------------------------------------------------------------------------
        for bi, blk in enumerate(p.blocks):
            bsum, bsq = sc.view(blk.o_sum, blk.ctot), sc.view(blk.o_sq, blk.ctot)
            for dl in blk.layers:
                s1, q1 = sc.view(dl.o_sum1, 128), sc.view(dl.o_sq1, 128)
                OC.conv_fprop_raw(dl.g1, blk.buf, self.w16(dl.name + "_1_conv/kernel"), dl.y1,
                                  pro_act=ACT_RELU, stat_sum=s1 if stats else None,
                                  stat_sq=q1 if stats else None,
                                  **self._bn_prologue(dl.bn0, training, bsum, bsq, blk.count))
                OC.conv_fprop_raw(dl.g2, dl.y1, self.w16(dl.name + "_2_conv/kernel"), blk.buf, y_coff=dl.cin,
                                  pro_act=ACT_RELU,
                                  stat_sum=bsum[dl.cin:dl.cin + GROWTH] if stats else None,
                                  stat_sq=bsq[dl.cin:dl.cin + GROWTH] if stats else None,
                                  **self._bn_prologue(dl.bn1, training, s1, q1, blk.count))
            if bi + 1 < len(p.blocks):
                nxt = p.blocks[bi + 1]
                OC.conv_fprop_raw(blk.gt, blk.buf, self.w16(blk.tconv), blk.t, pro_act=ACT_RELU,
                                  **self._bn_prologue(blk.bn_t, training, bsum, bsq, blk.count))
                K.pool_fwd(blk.t, nxt.buf, nxt.c0, 2, 2, 0, False)
                if stats:
                    K.channel_stats(nxt.buf, nxt.c0, sc.view(nxt.o_sum, nxt.c0), sc.view(nxt.o_sq, nxt.c0))
        # ---- head (final BN + ReLU folded into the pooling read) --------------------------------------
        lb = p.blocks[-1]
        self._bn_forward(lb.bn_t, training, sc.view(lb.o_sum, lb.ctot), sc.view(lb.o_sq, lb.ctot), lb.count)
        hl = p.head_layer
        wname, bname = hl.spec.params[0].name, hl.spec.params[1].name
        head_train = training and hl.index in self.trainable
        bwd = training and lb.needG
        if bwd:
            for blk in p.blocks:
                if blk.needG:
                    blk.G.zero_()
        K.head(lb.buf, lb.h * lb.w, lb.ctot, net.n_outputs, self.w32(wname), self.w32(bname), p.labels,
               p.logits, p.acc, pro_scale=lb.bn_t.scale, pro_shift=lb.bn_t.shift, pro_act=ACT_RELU,
               loss_kind=self.loss_kinds[net.loss.kind], acc_on_prob=net.acc_on_prob,
               threshold=net.acc_threshold, training=training, grad_scale=grad_scale,
               dw=self.gw(wname) if head_train else None, db=self.gw(bname) if head_train else None,
               gacc=lb.G if bwd else None, stat_a=lb.bn_t.sa if bwd else None,
               stat_b=lb.bn_t.sb if bwd else None)
        if not bwd:
            return
        self._backward(p)

    # -------------------------------------------------------------------------------------------------
    def _consumer_bn_backward(self, blk, bn: BNState) -> None:
        """A BN reading the block buffer has produced sa/sb in its consumer's dgrad epilogue. Its parameter
        gradients and its share of the batch-statistics correction (S1/S2) are NOT given a launch of their own:
        the next slice finalisation of the same block folds them in (csrc/elementwise.cu:idcb_bn_bwd_slice);
        whatever is still pending when the block's gradients are announced is flushed by _flush_pending."""
        if not bn.trainable:
            return
        self._flush_pending()
        self._pending = (blk, bn)

    _pending = None

    def _flush_pending(self) -> None:
        if self._pending is None:
            return
        blk, bn = self._pending
        self._pending = None
        sc = blk.scratch
        K.bn_bwd_coeffs(bn.sa, bn.sb, bn.mean, bn.rstd, bn.scale, blk.count, dgamma=bn.dgamma,
                        dbeta=bn.dbeta, acc_s1=sc.view(blk.o_s1, bn.c), acc_s2=sc.view(blk.o_s2, bn.c))

    def _finalize_slice(self, blk, c_lo: int, c_n: int, out: torch.Tensor) -> None:
        """out (bf16, c_n channels) = d(loss)/d(buffer[:, c_lo:c_lo+c_n]) including the BN-statistics
        terms of every consumer."""
        sc = blk.scratch
        pend = None
        if self._pending is not None:
            pblk, pbn = self._pending
            if pblk is blk and c_lo + c_n <= pbn.c:
                pend = (pbn.sa, pbn.sb, pbn.mean, pbn.rstd, pbn.scale, pbn.dgamma, pbn.dbeta)
                self._pending = None
            else:
                self._flush_pending()
        bn = blk.bn_t
        K.bn_bwd_slice(blk.G, blk.buf, out, c_lo, c_n, blk.count, sc.view(blk.o_s1, blk.ctot),
                       sc.view(blk.o_s2, blk.ctot), bn.mean, bn.rstd, pending=pend)

    def grad_ready_offsets(self) -> list:
        """after dense block bi's backward, every gradient at or above the block's first parameter is final"""
        net = self.net
        st = net.store
        offs = []
        bi = 1
        while True:
            names = [pp.name for l in net.layers if l.spec.name.startswith(f"conv{bi + 1}_block1_") for pp in l.spec.params]
            if not names:
                break
            offs.append(min(st.slots[n].offset for n in names))
            bi += 1
        self._block_offsets = offs            # index = dense block number (0-based)
        return list(reversed(offs))

    # ---- weight-gradient kernels run beside the data-gradient chain -----------------------------------
    # A layer's wgrad and its dgrad both start from the same dy and neither feeds the other; the layers are
    # small (a dgrad or wgrad of blocks 2-4 occupies a fraction of the SMs for a few microseconds), so the wgrad
    # is launched on a side stream and the critical path keeps only the dgrad chain. The fork/join edges are
    # ordinary stream waits, so they are captured into the step's CUDA graph as parallel branches.
    # Joins: before a buffer a pending wgrad reads (the block's dy3 / dy1 scratch) is overwritten — i.e. at the
    # start of the next layer — and before the block's gradients are announced as final.
    wgrad_side_stream = True
    # finish BN1's backward inside the 3x3 dgrad launch (IDCB_BN_TAIL=0: separate bn_bwd_apply launch)
    bn_tail = os.environ.get("IDCB_BN_TAIL", "1") != "0"
    # produce the next-lower layer's slice gradient inside the 1x1 dgrad launch (IDCB_SLICE_TAIL=0: own launch)
    slice_tail = os.environ.get("IDCB_SLICE_TAIL", "1") != "0"

    def _wgrad(self, *args, **kw) -> None:
        if not self.wgrad_side_stream:
            OC.conv_wgrad_raw(*args, **kw)
            return
        if getattr(self, "_wg_stream", None) is None:
            self._wg_stream = torch.cuda.Stream(device=self.net.device)
        main = torch.cuda.current_stream()
        self._wg_stream.wait_stream(main)
        with torch.cuda.stream(self._wg_stream):
            OC.conv_wgrad_raw(*args, **kw)
        self._wg_pending = True

    def _wgrad_join(self) -> None:
        if getattr(self, "_wg_pending", False):
            torch.cuda.current_stream().wait_stream(self._wg_stream)
            self._wg_pending = False

    def _backward(self, p) -> None:
        self._pending = None
        try:
            self._backward_impl(p)
            self._flush_pending()
        finally:
            self._wgrad_join()      # whichever way the truncated backward ended, the side stream rejoins

    def _backward_impl(self, p) -> None:
        if not hasattr(self, "_block_offsets"):
            self.grad_ready_offsets()
        ft = self.first_trainable
        tr = self.trainable
        sc = p.scratch
        nb = len(p.blocks)
        for bi in range(nb - 1, -1, -1):
            blk = p.blocks[bi]
            if not blk.bwd:
                break
            blk.scratch = sc
            if bi == nb - 1:
                # head already wrote G and the final BN's sums
                self._consumer_bn_backward(blk, blk.bn_t)
            else:
                nxt = p.blocks[bi + 1]
                # gradient of the transition's pooled output = finalized slice [0, c0) of the next block
                if ft <= blk.i_tconv:
                    self._finalize_slice(nxt, 0, nxt.c0, blk.dpool)
                    K.pool_bwd(blk.t, nxt.buf, blk.dpool, blk.dt, nxt.c0, 2, 2, 0, False)
                    if blk.i_tconv in tr:
                        self._wgrad(blk.gt, blk.buf, blk.dt, self.gw(blk.tconv), pro_scale=blk.bn_t.scale,
                                          pro_shift=blk.bn_t.shift, pro_act=ACT_RELU)
                    if ft <= blk.i_tbn:
                        OC.conv_dgrad_raw(blk.gt, blk.dt, self.w16(blk.tconv), None, dx_pitch=blk.ctot, mode=2,
                                          aux=blk.buf, bwd_scale=blk.bn_t.scale, bwd_shift=blk.bn_t.shift,
                                          bwd_act=ACT_RELU, gacc=blk.G, stat_a=blk.bn_t.sa, stat_b=blk.bn_t.sb)
                        self._consumer_bn_backward(blk, blk.bn_t)
            if not blk.needG:
                break
            slice_done = False      # this layer's slice gradient already came out of the 1x1 dgrad of the layer above
            for dl in reversed(blk.layers):
                if not dl.bwd:
                    break
                self._wgrad_join()          # the previous layer's wgrads still read dy1 (and the other dy3 copy)
                dy3 = blk.dy3[dl.li & 1]
                if not slice_done:
                    self._finalize_slice(blk, dl.cin, GROWTH, dy3)
                slice_done = False
                w2 = dl.name + "_2_conv/kernel"
                if dl.i_c2 in tr:
                    self._wgrad(dl.g2, dl.y1, dy3, self.gw(w2), pro_scale=dl.bn1.scale,
                                      pro_shift=dl.bn1.shift, pro_act=ACT_RELU)
                if ft > dl.i_bn1:
                    continue
                bn1 = dl.bn1
                # BN1's backward rides in the dgrad launch: column sums, a grid barrier with the tile parked in
                # TMEM, then dy1 straight from the accumulators (dpre1 never goes to memory). The native side
                # falls back to dgrad + bn_bwd_apply through dpre1 when the grid has more tiles than SMs (block 1).
                tail = None
                if self.bn_tail and ft <= dl.i_c1:
                    tail = OC.bn_tail(blk.dy1, bn1.mean, bn1.rstd, blk.count, counter=sc.view(dl.o_ctr, 1),
                                      dgamma=bn1.dgamma, dbeta=bn1.dbeta, frozen=not bn1.trainable)
                OC.conv_dgrad_raw(dl.g2, dy3, self.w16(w2), blk.dpre1, mode=2, aux=dl.y1,
                                  bwd_scale=bn1.scale, bwd_shift=bn1.shift, bwd_act=ACT_RELU,
                                  stat_a=bn1.sa, stat_b=bn1.sb, tail=tail)
                if ft > dl.i_c1:
                    # only BN1's own parameters are trainable below this point
                    if bn1.trainable:
                        K.bn_bwd_coeffs(bn1.sa, bn1.sb, bn1.mean, bn1.rstd, bn1.scale, blk.count,
                                        dgamma=bn1.dgamma, dbeta=bn1.dbeta)
                    continue
                if tail is None:
                    K.bn_bwd_apply(0 if bn1.trainable else 1, blk.dpre1, dl.y1, blk.dy1, 4 * GROWTH, blk.count,
                                   sa=bn1.sa, sb=bn1.sb, mean=bn1.mean, rstd=bn1.rstd, scale=bn1.scale,
                                   dgamma=bn1.dgamma, dbeta=bn1.dbeta)
                w1 = dl.name + "_1_conv/kernel"
                bn0 = dl.bn0
                if dl.i_c1 in tr:
                    self._wgrad(dl.g1, blk.buf, blk.dy1, self.gw(w1), pro_scale=bn0.scale,
                                      pro_shift=bn0.shift, pro_act=ACT_RELU)
                if ft > dl.i_bn0:
                    continue
                # The layer below produced buffer channels [cin - 32, cin): this dgrad is their last writer, so their
                # finished gradient (and this BatchNorm's bookkeeping) comes out of the same launch.
                below = blk.layers[dl.li - 1] if dl.li > 0 else None
                st = None
                if (self.slice_tail and below is not None and below.bwd and bn0.trainable
                        and OC.slice_tail_fits(blk.buf.numel() // blk.ctot, dl.cin)):
                    self._flush_pending()
                    st = OC.slice_tail(blk.dy3[below.li & 1], below.cin, sc.view(blk.o_s1, blk.ctot),
                                       sc.view(blk.o_s2, blk.ctot), blk.bn_t.mean, blk.bn_t.rstd, bn0.mean, bn0.rstd,
                                       blk.count, sc.view(dl.o_ctr2, 1), dgamma=bn0.dgamma, dbeta=bn0.dbeta)
                OC.conv_dgrad_raw(dl.g1, blk.dy1, self.w16(w1), None, dx_pitch=blk.ctot, mode=2, aux=blk.buf,
                                  bwd_scale=bn0.scale, bwd_shift=bn0.shift, bwd_act=ACT_RELU, gacc=blk.G,
                                  stat_a=bn0.sa, stat_b=bn0.sb, slice_tail=st)
                if st is not None:
                    slice_done = True
                else:
                    self._consumer_bn_backward(blk, bn0)
            self._wgrad_join()
            self._flush_pending()       # the block's first BatchNorm: its gradients are part of the announcement
            self.grads_ready(self._block_offsets[bi])
        # ---- stem ---------------------------------------------------------------------------------------
        if not p.stem_bwd:
            return
        b0 = p.blocks[0]
        self._finalize_slice(b0, 0, 64, p.dpool1)
        bn = p.bn_stem
        K.pool_bwd(p.s0, b0.buf, p.dpool1, p.dpre_stem, 64, 3, 2, 1, True, scale=bn.scale, shift=bn.shift,
                   act=ACT_RELU, stat_a=bn.sa, stat_b=bn.sb, argmax=p.pool1_arg)
        if ft > self.net.get_layer("conv1/conv").index:
            if bn.trainable:
                K.bn_bwd_coeffs(bn.sa, bn.sb, bn.mean, bn.rstd, bn.scale, float(p.s0.numel() // 64),
                                dgamma=bn.dgamma, dbeta=bn.dbeta)
            return
        K.bn_bwd_apply(0 if bn.trainable else 1, p.dpre_stem, p.s0, p.ds0, 64, float(p.s0.numel() // 64),
                       sa=bn.sa, sb=bn.sb, mean=bn.mean, rstd=bn.rstd, scale=bn.scale,
                       dgamma=bn.dgamma, dbeta=bn.dbeta)
        OC.conv_wgrad_raw(p.g_stem, p.x16, p.ds0, self.gw("conv1/conv/kernel"))
