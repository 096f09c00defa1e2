"""Losses used by the reference scripts, stated on logits.

* ``BinaryCrossentropy(from_logits=True)``  — dist_model_tf_mobile.py:131, dist_model_tf_vgg.py:131,
  fed_model.py:127,203, secure_fed_model.py:96:   mean_b( max(x,0) - x*z + log1p(exp(-|x|)) )
* ``CategoricalCrossentropy(from_logits=True)`` — dist_model_tf_dense.py:143. The reference feeds it
  integer CIFAR labels, which is a defect; the intent is 10-way softmax CE, so integer
  labels are accepted (sparse) as well as one-hot rows.

``__call__`` returns the per-example loss vector; reduction (sum / global batch) is done by the
training engine so that data-parallel replicas scale by the GLOBAL batch size like Keras does under a
``tf.distribute`` strategy.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class Loss:
    name = "loss"
    kind = "none"

    def per_example(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def __call__(self, logits, labels):
        return self.per_example(logits, labels).mean()


class BinaryCrossentropy(Loss):
    kind = "bce"

    def __init__(self, from_logits: bool = True):
        if not from_logits:
            raise NotImplementedError("only from_logits=True is used by the reference scripts")

    def per_example(self, logits, labels):
        x = logits.reshape(logits.shape[0], -1).float()
        z = labels.reshape(labels.shape[0], -1).to(x.dtype)
        l = torch.clamp(x, min=0) - x * z + torch.log1p(torch.exp(-x.abs()))
        return l.mean(dim=1)


class CategoricalCrossentropy(Loss):
    kind = "cce"

    def __init__(self, from_logits: bool = True):
        if not from_logits:
            raise NotImplementedError("only from_logits=True is used by the reference scripts")

    def per_example(self, logits, labels):
        x = logits.float()
        if labels.dim() == x.dim() and labels.shape[-1] == x.shape[-1] and labels.dtype.is_floating_point:
            return -(labels * F.log_softmax(x, dim=-1)).sum(-1)
        return F.cross_entropy(x, labels.reshape(-1).long(), reduction="none")


SparseCategoricalCrossentropy = CategoricalCrossentropy


def get(loss) -> Loss:
    if isinstance(loss, Loss):
        return loss
    table = {"binary_crossentropy": BinaryCrossentropy, "bce": BinaryCrossentropy,
             "categorical_crossentropy": CategoricalCrossentropy, "cce": CategoricalCrossentropy,
             "sparse_categorical_crossentropy": CategoricalCrossentropy}
    if isinstance(loss, str) and loss in table:
        return table[loss](from_logits=True)
    raise ValueError(f"unknown loss {loss!r}")
