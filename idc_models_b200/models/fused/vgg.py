"""Fused sm_100a pipeline for chain CNNs without normalisation: VGG16 + GAP + Dense head
(/root/reference/dist_model_tf_vgg.py:119-129, fed_model.py:113-123).

forward : 13 x [tcgen05 3x3 conv, bias + ReLU in the epilogue], 5 x 2x2 max-pool, one head kernel.
backward: only down to the first trainable layer (``fine_tune_at``): per conv one wgrad and one dgrad
          whose epilogue applies the previous layer's ReLU mask and accumulates its bias gradient;
          max-pool backward does the same for the conv in front of it. No cuDNN, no cuBLAS.
"""
from __future__ import annotations

import torch

from ...ops import conv as OC
from ...ops import kernels as K
from .common import ACT_NONE, ACT_RELU, FusedExecutor, Scratch


class _Plan:
    pass


class ChainExecutor(FusedExecutor):
    """conv(+bias,+relu) / maxpool chains followed by gap+dense."""

    def build_plan(self, batch: int):
        net = self.net
        dev = net.device
        p = _Plan()
        p.batch = batch
        h, w, c = net.input_shape
        p.x32 = torch.zeros(batch, h, w, c, dtype=torch.float32, device=dev)
        p.labels = torch.zeros(batch, dtype=torch.float32, device=dev)
        p.acc = torch.zeros(4, dtype=torch.float32, device=dev)
        p.logits = torch.zeros(batch, net.n_outputs, dtype=torch.float32, device=dev)
        cpad = -(-c // 8) * 8
        p.x16 = torch.zeros(batch, h, w, cpad, dtype=torch.bfloat16, device=dev)
        p.steps = []          # forward ops in order
        cur, cur_shape = p.x16, (h, w, cpad)
        for layer in net.layers:
            sp = layer.spec
            if sp.kind == "input":
                continue
            if sp.kind == "conv":
                hh, ww, cin = cur_shape
                g = OC.geom(hh, ww, cin, sp.attrs["filters"] if "filters" in sp.attrs else sp.out_shape[-1],
                            sp.attrs["k"], sp.attrs["stride"], sp.attrs["padding"])
                y = torch.zeros(batch, g.ho, g.wo, g.cout, dtype=torch.bfloat16, device=dev)
                st = dict(kind="conv", layer=layer, geom=g, x=cur, y=y, dy=None,
                          relu=sp.attrs.get("activation") == "relu", bias=sp.attrs["use_bias"])
                p.steps.append(st)
                cur, cur_shape = y, (g.ho, g.wo, g.cout)
            elif sp.kind == "maxpool":
                hh, ww, cc = cur_shape
                k, s = sp.attrs["k"], sp.attrs["stride"]
                ho, wo = (hh - k) // s + 1, (ww - k) // s + 1
                y = torch.zeros(batch, ho, wo, cc, dtype=torch.bfloat16, device=dev)
                # one byte per output element: which window element won (the backward pass reads it instead of
                # re-comparing values)
                arg = torch.zeros(batch, ho, wo, cc, dtype=torch.uint8, device=dev)
                p.steps.append(dict(kind="maxpool", layer=layer, x=cur, y=y, k=k, s=s, c=cc, dy=None, arg=arg))
                cur, cur_shape = y, (ho, wo, cc)
            elif sp.kind == "gap":
                p.feat, p.feat_shape = cur, cur_shape
            elif sp.kind == "dense":
                p.head_layer = layer
            else:
                raise NotImplementedError(f"ChainExecutor: layer kind {sp.kind}")
        # gradient buffers, only where backward will run
        for st in p.steps:
            if self.needs_grad_below(st["layer"].index):
                st["dy"] = torch.zeros_like(st["y"])
        return p

    wgrad_side_stream = True

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

    def _side_join(self) -> None:
        if getattr(self, "_wg_pending", False):
            torch.cuda.current_stream().wait_stream(self._wg_stream)
            self._wg_pending = False

    def grad_ready_offsets(self) -> list:
        st = self.net.store
        offs = [st.slots[l.spec.params[0].name].offset for l in self.net.layers
                if l.spec.kind == "conv" and l.index in self.trainable]
        return sorted(offs, reverse=True)

    def run(self, p, training: bool, grad_scale: float) -> None:
        net = self.net
        p.acc.zero_()
        K.pack_input(p.x32, p.x16)
        for st in p.steps:
            if st["kind"] == "conv":
                name = st["layer"].spec.params[0].name
                bias = self.w32(st["layer"].spec.params[1].name) if st["bias"] else None
                OC.conv_fprop_raw(st["geom"], st["x"], self.w16(name), st["y"], epi_shift=bias,
                                  epi_act=ACT_RELU if st["relu"] else ACT_NONE)
            else:
                K.pool_fwd(st["x"], st["y"], st["c"], st["k"], st["s"], 0, True, argmax=st["arg"] if training else None)
        hl = p.head_layer
        hh, ww, cc = p.feat_shape
        wname, bname = hl.spec.params[0].name, hl.spec.params[1].name
        head_train = training and hl.index in self.trainable
        last = p.steps[-1]
        need_dx = training and last["dy"] is not None
        K.head(p.feat, hh * ww, cc, net.n_outputs, self.w32(wname), self.w32(bname), p.labels, p.logits,
               p.acc, loss_kind=self.loss_kinds[net.loss.kind], acc_on_prob=net.acc_on_prob,
               threshold=net.acc_threshold, training=training, grad_scale=grad_scale,
               dw=self.gw(wname) if head_train else None, db=self.gw(bname) if head_train else None,
               dx=last["dy"] if need_dx else None)
        if not training:
            return
        # ---- backward through the chain -----------------------------------------------------------
        # invariant: st["dy"] holds dL/d(output of st) ; for a ReLU conv it is turned into dpre (masked)
        # by whoever writes it (the consumer's dgrad epilogue / pool backward), except for the last
        # step, whose dy comes from the head and is masked here if needed.
        n = len(p.steps)
        if need_dx and last["kind"] == "conv" and last["relu"]:
            raise NotImplementedError("chain must end with a pooling layer before the head")
        for i in range(n - 1, -1, -1):
            st = p.steps[i]
            if st["dy"] is None:
                break
            prev = p.steps[i - 1] if i > 0 else None
            prev_needs = prev is not None and prev["dy"] is not None
            if st["kind"] == "maxpool":
                # producer of st["x"] is a ReLU conv: fuse its ReLU mask and bias gradient
                if not prev_needs:
                    continue
                pl = prev["layer"]
                db = None
                if prev["kind"] == "conv" and prev["bias"] and pl.index in self.trainable:
                    db = self.gw(pl.spec.params[1].name)
                K.pool_bwd(st["x"], st["y"], st["dy"], prev["dy"], st["c"], st["k"], st["s"], 0, True,
                           relu_mask=(prev["kind"] == "conv" and prev["relu"]), stat_a=db, argmax=st["arg"])
            else:
                layer = st["layer"]
                g = st["geom"]
                wname = layer.spec.params[0].name
                if layer.index in self.trainable:
                    # beside the data-gradient kernel of the same layer (both only read x / dy; see DenseNet executor)
                    self._side(OC.conv_wgrad_raw, g, st["x"], st["dy"], self.gw(wname))
                if prev_needs:
                    pl = prev["layer"]
                    if prev["kind"] == "conv":
                        db = self.gw(pl.spec.params[1].name) if (prev["bias"] and pl.index in self.trainable) else None
                        OC.conv_dgrad_raw(g, st["dy"], self.w16(wname), prev["dy"],
                                          mode=1 if prev["relu"] else 0, aux=st["x"] if prev["relu"] else None,
                                          stat_a=db)
                    else:
                        OC.conv_dgrad_raw(g, st["dy"], self.w16(wname), prev["dy"], mode=0)
                # this layer's weight gradient and (written by its consumer earlier) bias gradient are final
                if self.grad_hook is not None:
                    self._side_join()
                self.grads_ready(self.net.store.slots[wname].offset)
        self._side_join()
