"""Streaming metrics with Keras semantics.

Quirk kept on purpose (SURVEY.md 2.A "Known defects"): ``metrics=['accuracy']`` with a 1-unit logit
output resolves to Keras ``binary_accuracy``, which thresholds the RAW LOGIT at 0.5 (not the sigmoid),
and ``BinaryAccuracy()`` in the FL scripts does the same (dist_model_tf_mobile.py:132, fed_model.py:128,
secure_fed_model.py:97). Pass ``threshold_on='prob'`` for the conventional sigmoid>=0.5 rule.

``auroc`` mirrors secure_fed_model.py:81-82: sklearn ``roc_auc_score`` evaluated per batch and averaged
by the metric container (batches with a single class are skipped instead of raising).

Every metric accumulates ``(total, count)`` so that replicas / federated clients can be combined by
summing accumulators.
"""
from __future__ import annotations

import numpy as np
import torch


class Metric:
    name = "metric"

    def __init__(self):
        self.total = 0.0
        self.count = 0.0

    def reset(self):
        self.total = 0.0
        self.count = 0.0

    def batch_stats(self, logits: torch.Tensor, labels: torch.Tensor) -> tuple[float, float]:
        raise NotImplementedError

    def update(self, logits, labels):
        t, c = self.batch_stats(logits, labels)
        self.total += float(t)
        self.count += float(c)

    def merge(self, total: float, count: float):
        self.total += total
        self.count += count

    def result(self) -> float:
        return self.total / self.count if self.count else 0.0


class BinaryAccuracy(Metric):
    def __init__(self, name="binary_accuracy", threshold=0.5, threshold_on="logit"):
        super().__init__()
        self.name = name
        self.threshold = threshold
        self.threshold_on = threshold_on

    def batch_stats(self, logits, labels):
        x = logits.reshape(-1).float()
        if self.threshold_on == "prob":
            x = torch.sigmoid(x)
        pred = (x > self.threshold).to(torch.float32)
        z = labels.reshape(-1).to(torch.float32)
        return (pred == z).sum().item(), z.numel()


class CategoricalAccuracy(Metric):
    def __init__(self, name="accuracy"):
        super().__init__()
        self.name = name

    def batch_stats(self, logits, labels):
        pred = logits.argmax(-1)
        if labels.dim() == logits.dim() and labels.shape[-1] == logits.shape[-1]:
            labels = labels.argmax(-1)
        z = labels.reshape(-1).long()
        return (pred == z).sum().item(), z.numel()


def roc_auc_score(y_true: np.ndarray, y_score: np.ndarray) -> float:
    """Rank-based AUROC (ties get average rank) — same value as sklearn.metrics.roc_auc_score."""
    y_true = np.asarray(y_true).reshape(-1).astype(bool)
    y_score = np.asarray(y_score, dtype=np.float64).reshape(-1)
    n_pos = int(y_true.sum())
    n_neg = y_true.size - n_pos
    if n_pos == 0 or n_neg == 0:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    order = np.argsort(y_score, kind="mergesort")
    s = y_score[order]
    ranks = np.empty(s.size, dtype=np.float64)
    i = 0
    while i < s.size:
        j = i
        while j + 1 < s.size and s[j + 1] == s[i]:
            j += 1
        ranks[i:j + 1] = 0.5 * (i + j) + 1.0
        i = j + 1
    r = np.empty_like(ranks)
    r[order] = ranks
    return float((r[y_true].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


class AUROC(Metric):
    """Per-batch AUROC averaged over batches (the py_func metric of secure_fed_model.py:81-82)."""

    def __init__(self, name="auroc"):
        super().__init__()
        self.name = name

    def batch_stats(self, logits, labels):
        y = labels.reshape(-1).detach().cpu().numpy()
        s = logits.reshape(-1).detach().float().cpu().numpy()
        try:
            return roc_auc_score(y, s), 1.0
        except ValueError:
            return 0.0, 0.0


def auroc(y_true=None, y_pred=None):
    """Callable alias so scripts can write ``metrics=[BinaryAccuracy(), auroc]`` like the reference."""
    if y_true is None:
        return AUROC()
    return roc_auc_score(np.asarray(y_true), np.asarray(y_pred))


class FunctionMetric(Metric):
    """A plain ``fn(y_true, y_pred)`` passed in ``metrics=[...]`` (Keras calls it per batch and averages the values
    over the batches; the reference's own ``auroc`` py_func wrapper is one, secure_fed_model.py:81-82,97). A batch
    on which the function raises ``ValueError`` (AUROC of a single-class batch) is left out of the average."""

    def __init__(self, fn):
        super().__init__()
        self.fn = fn
        self.name = getattr(fn, "__name__", "metric")

    def batch_stats(self, logits, labels):
        y = labels.detach().cpu().numpy().reshape(logits.shape[0], -1)
        s = logits.detach().float().cpu().numpy().reshape(logits.shape[0], -1)
        try:
            return float(np.mean(np.asarray(self.fn(y, s), dtype=np.float64))), 1.0
        except ValueError:
            return 0.0, 0.0


def get(m, n_outputs: int) -> Metric:
    if isinstance(m, Metric):
        return m
    if m is auroc:
        return AUROC()
    if isinstance(m, str):
        if m in ("accuracy", "acc"):
            return BinaryAccuracy(name="accuracy") if n_outputs == 1 else CategoricalAccuracy(name="accuracy")
        if m in ("binary_accuracy",):
            return BinaryAccuracy()
        if m in ("auroc", "auc"):
            return AUROC()
    if callable(m):
        import inspect
        try:
            required = [p for p in inspect.signature(m).parameters.values()
                        if p.default is p.empty and p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        except (TypeError, ValueError):
            required = []
        if not required:                       # a metric class / factory
            r = m()
            if isinstance(r, Metric):
                return r
        elif len(required) == 2:               # fn(y_true, y_pred)
            return FunctionMetric(m)
    raise ValueError(f"unknown metric {m!r}")
