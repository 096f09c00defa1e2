"""Plain-PyTorch fp32 executor for a :class:`ModelSpec` (runs anywhere; it IS the CPU path and the
numerical oracle for the fused sm_100a pipelines).

Semantics follow TF2 Keras as used by the reference scripts: a BatchNorm layer
whose ``trainable`` flag is False runs in inference mode even during ``fit``; a trainable one uses
per-replica batch statistics while training and updates its moving statistics with the layer's
momentum; ``evaluate``/``predict`` always use moving statistics.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class RefExecutor:
    def __init__(self, net, emulate_bf16: bool = False):
        self.net = net
        self._leaves: dict[str, torch.Tensor] = {}
        self._leaf_dev = None
        # Oracle mode for the fused bf16 pipelines: every layer output is rounded to bf16 (straight-
        # through for autograd), which is where the fused kernels round when they store activations.
        # All arithmetic stays fp32. Without it, discontinuous ops (max-pool argmax, ReLU masks, tiny-
        # batch BatchNorm) amplify the harmless storage rounding into large apparent differences.
        self.emulate_bf16 = emulate_bf16

    # parameters as autograd leaves whose .grad aliases the flat gradient buffer
    def _leaf(self, name: str, requires_grad: bool) -> torch.Tensor:
        st = self.net.store
        if self._leaf_dev != (st.master.data_ptr(), st.grad.data_ptr()):
            self._leaves.clear()
            self._leaf_dev = (st.master.data_ptr(), st.grad.data_ptr())
        t = self._leaves.get(name)
        if t is None:
            t = st.view(name)
            if st.slots[name].trainable_kind:
                t = t.detach()
                t.requires_grad_(True)
                t.grad = st.grad_view(name)
            self._leaves[name] = t
        if st.slots[name].trainable_kind:
            t.requires_grad_(requires_grad)
            if requires_grad and t.grad is None:
                t.grad = st.grad_view(name)
        return t

    def forward(self, x: torch.Tensor, training: bool, upto: str | None = None,
                generator: torch.Generator | None = None, capture: dict | None = None) -> torch.Tensor:
        """x: [B, H, W, C] float32 in [0,1]. Returns the output of the last layer (or of ``upto``).
        ``capture``: filled with every layer's output (NCHW, gradients retained) for layer-isolated tests."""
        net = self.net
        vals: dict[str, torch.Tensor] = {}
        out = None
        for li, layer in enumerate(net.layers):
            sp = layer.spec
            k = sp.kind
            trainable = layer.trainable and training
            ins = [vals[n] for n in sp.inputs]
            if k == "input":
                out = x.permute(0, 3, 1, 2).to(torch.float32)
            elif k == "zeropad":
                (pt, pb), (pl, pr) = sp.attrs["pad"]
                out = F.pad(ins[0], (pl, pr, pt, pb))
            elif k == "conv":
                w = self._leaf(sp.params[0].name, trainable).permute(0, 3, 1, 2)
                b = self._leaf(sp.params[1].name, trainable) if sp.attrs["use_bias"] else None
                xin = ins[0]
                st, kk = sp.attrs["stride"], sp.attrs["k"]
                if sp.attrs["padding"] == "same":
                    from ..ops.conv import same_padding
                    (pt, pb), (pl, pr) = same_padding(xin.shape[2], kk, st), same_padding(xin.shape[3], kk, st)
                    xin = F.pad(xin, (pl, pr, pt, pb))
                out = F.conv2d(xin, w, b, stride=st)
                if sp.attrs.get("activation") == "relu":
                    out = torch.relu(out)
            elif k == "dwconv":
                w = self._leaf(sp.params[0].name, trainable)  # [kh, kw, C]
                c = w.shape[-1]
                xin = ins[0]
                st, kk = sp.attrs["stride"], sp.attrs["k"]
                if sp.attrs["padding"] == "same":
                    from ..ops.conv import same_padding
                    (pt, pb), (pl, pr) = same_padding(xin.shape[2], kk, st), same_padding(xin.shape[3], kk, st)
                    xin = F.pad(xin, (pl, pr, pt, pb))
                out = F.conv2d(xin, w.permute(2, 0, 1).unsqueeze(1), None, stride=st, groups=c)
            elif k == "bn":
                g = self._leaf(sp.params[0].name, trainable)
                bta = self._leaf(sp.params[1].name, trainable)
                mean = self._leaf(sp.params[2].name, False)
                var = self._leaf(sp.params[3].name, False)
                xin = ins[0]
                if trainable:
                    # batch statistics; moving stats updated in place (Keras momentum convention)
                    out = F.batch_norm(xin, None, None, g, bta, True, 0.0, sp.attrs["eps"])
                    with torch.no_grad():
                        dims = (0, 2, 3)
                        n = xin.numel() // xin.shape[1]
                        bm = xin.mean(dims)
                        bv = xin.var(dims, unbiased=False) * (n / max(n - 1, 1))
                        mom = sp.attrs["momentum"]
                        mean.mul_(mom).add_(bm, alpha=1 - mom)
                        var.mul_(mom).add_(bv, alpha=1 - mom)
                else:
                    out = F.batch_norm(xin, mean, var, g, bta, False, 0.0, sp.attrs["eps"])
            elif k == "act":
                mv = sp.attrs.get("max_value")
                out = torch.relu(ins[0]) if mv is None else torch.clamp(ins[0], 0.0, float(mv))
            elif k == "maxpool":
                out = F.max_pool2d(ins[0], sp.attrs["k"], sp.attrs["stride"])
            elif k == "avgpool":
                out = F.avg_pool2d(ins[0], sp.attrs["k"], sp.attrs["stride"])
            elif k == "concat":
                out = torch.cat(ins, dim=1)
            elif k == "add":
                out = ins[0] + ins[1]
            elif k == "gap":
                out = ins[0].mean(dim=(2, 3))
            elif k == "flatten":
                out = ins[0].permute(0, 2, 3, 1).reshape(ins[0].shape[0], -1)  # Keras flattens NHWC
            elif k == "dropout":
                if training:
                    p = sp.attrs["rate"]
                    mask = (torch.rand(ins[0].shape, generator=generator, device="cpu") >= p).to(ins[0])
                    out = ins[0] * mask / (1 - p)
                else:
                    out = ins[0]
            elif k == "dense":
                w = self._leaf(sp.params[0].name, trainable)
                b = self._leaf(sp.params[1].name, trainable)
                out = F.linear(ins[0], w, b)
                if sp.attrs.get("activation") == "relu":
                    out = torch.relu(out)
            else:  # pragma: no cover
                raise NotImplementedError(k)
            if self.emulate_bf16 and k in ("input", "conv", "dwconv", "bn", "act", "maxpool", "avgpool", "add"):
                out = out + (out.to(torch.bfloat16).to(out.dtype) - out).detach()
            vals[sp.name] = out
            if capture is not None:
                if out.requires_grad:
                    out.retain_grad()
                capture[sp.name] = out
            if upto is not None and sp.name == upto:
                break
        return out

    # ---- executor protocol (shared with the fused CUDA executors) --------------------------------
    def prepare(self) -> None:
        pass

    def _stats(self, logits, y, per_ex):
        net = self.net
        with torch.no_grad():
            if net.n_outputs == 1:
                x = logits.reshape(-1)
                if net.acc_on_prob:
                    x = torch.sigmoid(x)
                correct = ((x > net.acc_threshold).float() == y.reshape(-1).float()).sum()
            else:
                correct = (logits.argmax(-1) == y.reshape(-1).long()).sum()
            return torch.stack([per_ex.sum().float(), correct.float(),
                                torch.tensor(float(y.shape[0]), device=logits.device)])

    def train_batch(self, x: torch.Tensor, y: torch.Tensor, global_batch: int):
        """Forward + backward for this replica's slice; gradients of (sum_i loss_i) / global_batch
        are accumulated into the flat gradient buffer. Returns ([loss_sum, correct, count], logits)."""
        net = self.net
        x = x.to(net.device, torch.float32)
        y = y.to(net.device)
        if x.shape[0] == 0:
            z = torch.zeros(3, device=net.device)
            return z, torch.zeros(0, net.n_outputs, device=net.device)
        logits = self.forward(x, training=True, generator=getattr(net, "_dropout_gen", None))
        per_ex = net.loss.per_example(logits, y)
        (per_ex.sum() / float(global_batch)).backward()
        return self._stats(logits.detach(), y, per_ex.detach()), logits.detach()

    def eval_batch(self, x: torch.Tensor, y: torch.Tensor):
        net = self.net
        x = x.to(net.device, torch.float32)
        y = y.to(net.device)
        with torch.no_grad():
            logits = self.forward(x, training=False)
            per_ex = net.loss.per_example(logits, y)
        return self._stats(logits, y, per_ex), logits

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            return self.forward(x.to(self.net.device, torch.float32), training=False)
