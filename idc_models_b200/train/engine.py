"""compile / fit / evaluate / predict with Keras semantics:

* ``fit`` runs epochs ``initial_epoch .. epochs-1`` — the reference's phase 2 passes
  ``initial_epoch=history.epoch[-1]`` (= 9) and ``epochs=20``, i.e. 11 epochs
  (/root/reference/dist_model_tf_mobile.py:157-159);
* after every epoch the FULL ``validation_data`` is evaluated; ``History.history`` has the keys
  ``loss / <metric> / val_loss / val_<metric>`` and ``History.epoch`` the epoch indices;
* ``evaluate(ds, steps=20)`` consumes at most 20 batches and returns ``[loss, metric...]``
  (dist_model_tf_mobile.py:134);
* under a strategy each replica takes its slice of every global batch, the loss is divided by the
  GLOBAL batch size, gradients are summed over replicas and the same update is applied everywhere
  (SURVEY.md B8, B12); loss / metric accumulators are summed over replicas (C2).

Recompiling creates fresh optimizer state (dist_model_tf_mobile.py:152-154).
"""
from __future__ import annotations

import sys
import time

import torch

from ..optim import optimizers as O
from ..utils.trace import JsonlLogger, nvtx_range
from . import losses as L
from . import metrics as M


class History:
    def __init__(self):
        self.history: dict[str, list] = {}
        self.epoch: list[int] = []

    def _append(self, epoch: int, logs: dict) -> None:
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(v)


class Callback:
    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None): ...
    def on_epoch_begin(self, epoch, logs=None): ...
    def on_epoch_end(self, epoch, logs=None): ...
    def on_train_end(self, logs=None): ...


class ModelCheckpoint(Callback):
    """``tf.keras.callbacks.ModelCheckpoint(filepath, save_weights_only=True, verbose=1)`` — weights
    are written at the end of every epoch, overwriting (fed_model.py:100-105,135)."""

    def __init__(self, filepath: str, save_weights_only: bool = True, verbose: int = 0, **kw):
        self.filepath, self.verbose = filepath, verbose

    def on_epoch_end(self, epoch, logs=None):
        path = self.filepath.format(epoch=epoch + 1, **(logs or {}))
        if self.verbose:
            print(f"\nEpoch {epoch + 1:05d}: saving model to {path}")
        self.model.save_weights(path)


class EarlyStopping(Callback):
    """``tf.keras.callbacks.EarlyStopping``: stop when ``monitor`` has not improved by ``min_delta`` for ``patience``
    epochs; optionally put the best weights back."""

    def __init__(self, monitor: str = "val_loss", min_delta: float = 0.0, patience: int = 0, mode: str = "auto",
                 restore_best_weights: bool = False, verbose: int = 0, **kw):
        self.monitor, self.min_delta, self.patience = monitor, abs(float(min_delta)), int(patience)
        self.restore_best_weights, self.verbose = restore_best_weights, verbose
        self.larger_is_better = mode == "max" or (mode == "auto" and ("acc" in monitor or "auc" in monitor))

    def on_train_begin(self, logs=None):
        self.best, self.wait, self.stopped_epoch, self.best_weights = None, 0, None, None

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        better = self.best is None or (cur > self.best + self.min_delta if self.larger_is_better
                                       else cur < self.best - self.min_delta)
        if better:
            self.best, self.wait = cur, 0
            if self.restore_best_weights:
                self.best_weights = self.model.get_weights()
            return
        self.wait += 1
        if self.wait >= self.patience:
            self.stopped_epoch = epoch
            self.model.stop_training = True
            if self.restore_best_weights and self.best_weights is not None:
                self.model.set_weights(self.best_weights)
            if self.verbose:
                print(f"Epoch {epoch + 1:05d}: early stopping")


class LearningRateScheduler(Callback):
    """``tf.keras.callbacks.LearningRateScheduler(schedule)``: ``schedule(epoch, lr) -> lr`` (or ``schedule(epoch)``)
    at the start of every epoch. The optimizer reads its learning rate when a step is launched, so no re-compile is
    needed (a data-parallel step whose all-reduce is captured in a CUDA graph re-captures once per new value)."""

    def __init__(self, schedule, verbose: int = 0):
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        opt = self.model.optimizer
        try:
            lr = self.schedule(epoch, opt.learning_rate)
        except TypeError:
            lr = self.schedule(epoch)
        opt.learning_rate = float(lr)
        if self.verbose:
            print(f"\nEpoch {epoch + 1:05d}: LearningRateScheduler setting learning rate to {opt.learning_rate}.")

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = self.model.optimizer.learning_rate


class ReduceLROnPlateau(Callback):
    """``tf.keras.callbacks.ReduceLROnPlateau``: multiply the learning rate by ``factor`` after ``patience`` epochs
    without improvement of ``monitor`` (then wait ``cooldown`` epochs), never below ``min_lr``."""

    def __init__(self, monitor: str = "val_loss", factor: float = 0.1, patience: int = 10, min_delta: float = 1e-4,
                 cooldown: int = 0, min_lr: float = 0.0, mode: str = "auto", verbose: int = 0, **kw):
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0")
        self.monitor, self.factor, self.patience, self.min_delta = monitor, float(factor), int(patience), abs(float(min_delta))
        self.cooldown, self.min_lr, self.verbose = int(cooldown), float(min_lr), verbose
        self.larger_is_better = mode == "max" or (mode == "auto" and ("acc" in monitor or "auc" in monitor))

    def on_train_begin(self, logs=None):
        self.best, self.wait, self.cooldown_counter = None, 0, 0

    def on_epoch_end(self, epoch, logs=None):
        opt = self.model.optimizer
        if logs is not None:
            logs["lr"] = opt.learning_rate
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        better = self.best is None or (cur > self.best + self.min_delta if self.larger_is_better
                                       else cur < self.best - self.min_delta)
        if better:
            self.best, self.wait = cur, 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience and opt.learning_rate > self.min_lr:
                opt.learning_rate = max(opt.learning_rate * self.factor, self.min_lr)
                self.cooldown_counter, self.wait = self.cooldown, 0
                if self.verbose:
                    print(f"\nEpoch {epoch + 1:05d}: ReduceLROnPlateau reducing learning rate to {opt.learning_rate}.")


class CSVLogger(Callback):
    """``tf.keras.callbacks.CSVLogger(filename, separator=',', append=False)``: one row per epoch."""

    def __init__(self, filename: str, separator: str = ",", append: bool = False):
        self.filename, self.sep, self.append = filename, separator, append
        self.keys = None

    def on_train_begin(self, logs=None):
        import os
        self._fh = open(self.filename, "a" if self.append else "w")
        self.keys = None
        self._header = not (self.append and os.path.getsize(self.filename) > 0)

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self.keys is None:
            self.keys = sorted(logs)
            if self._header:
                self._fh.write(self.sep.join(["epoch"] + self.keys) + "\n")
        self._fh.write(self.sep.join([str(epoch)] + [repr(float(logs.get(k, float("nan")))) for k in self.keys]) + "\n")
        self._fh.flush()

    def on_train_end(self, logs=None):
        self._fh.close()


def compile_model(net, optimizer, loss, metrics) -> None:
    net.loss = L.get(loss if loss is not None else ("bce" if net.n_outputs == 1 else "cce"))
    net.metrics = [M.get(m, net.n_outputs) for m in metrics]
    # accuracy rule used by the in-kernel counters (and by the reference executor)
    net.acc_threshold, net.acc_on_prob = 0.5, False
    for m in net.metrics:
        if isinstance(m, M.BinaryAccuracy):
            net.acc_threshold, net.acc_on_prob = m.threshold, m.threshold_on == "prob"
    net.strategy.prepare_store(net.store)
    opt = O.get(optimizer)
    if opt.store is not None:          # a bound optimizer object reused for another compile
        opt = opt.clone()
    ranges = net.store.ranges(net.trainable_layer_indices())
    net.optimizer = opt.bind(net.store, ranges)
    net.store.grad.zero_()
    net._dirty = False
    net.executor.prepare()
    # a strategy may hook into the executor's backward pass (gradient all-reduce overlapped with it)
    if hasattr(net.executor, "grad_hook"):
        net.executor.grad_hook = None
    after = getattr(net.strategy, "after_compile", None)
    if after is not None:
        after(net)


def _check_compiled(net):
    if net.optimizer is None or net.loss is None:
        raise RuntimeError("model.compile(...) must be called before fit/evaluate")
    if net._dirty:
        # Keras semantics: changing `trainable` after compile has no effect until the next compile.
        pass


def _builtin_metric(m) -> bool:
    return isinstance(m, (M.BinaryAccuracy, M.CategoricalAccuracy))


def _as_batches(data):
    if isinstance(data, tuple) and len(data) == 2 and isinstance(data[0], torch.Tensor):
        return [data]
    return data


class _Accum:
    """Device-side (loss_sum, correct, count) accumulator + host-side extra metrics."""

    def __init__(self, net):
        self.net = net
        self.acc = torch.zeros(3, dtype=torch.float32, device=net.device)
        self.gacc = torch.zeros(3, dtype=torch.float32, device=net.device)   # None once a step came without global sums
        self.extra = [m for m in net.metrics if not _builtin_metric(m)]
        for m in self.extra:
            m.reset()

    def add(self, stats, logits, y, global_stats=None):
        """``global_stats``: the same three slots already summed over replicas (they rode on the fused
        all-reduce kernel's barrier); when every step provides them no collective is needed at the end."""
        self.acc += stats.to(self.acc.dtype)
        if global_stats is not None and self.gacc is not None:
            self.gacc += global_stats.to(self.acc.dtype)
        else:
            self.gacc = None
        for m in self.extra:
            if logits is not None and logits.shape[0]:
                m.update(logits.float().cpu(), y.cpu())

    def result(self, prefix: str = "") -> dict:
        net = self.net
        acc = self.gacc if self.gacc is not None else net.strategy.reduce_scalars(self.acc.clone())
        loss_sum, correct, count = (float(v) for v in acc.tolist())
        out = {prefix + "loss": loss_sum / count if count else 0.0}
        for m in net.metrics:
            if _builtin_metric(m):
                out[prefix + m.name] = correct / count if count else 0.0
            else:
                t = torch.tensor([m.total, m.count], dtype=torch.float64)
                if net.strategy.world.is_distributed:
                    t = net.strategy.reduce_scalars(t.to(net.device)).cpu()
                out[prefix + m.name] = float(t[0] / t[1]) if t[1] else 0.0
        return out


def train_on_batch(net, x, y) -> dict:
    _check_compiled(net)
    n = x.shape[0]
    sl = net.strategy.local_slice(n)
    stats, logits = net.executor.train_batch(x[sl], y[sl], n)
    net.strategy.reduce_and_apply(net.store, net.optimizer, 1.0)
    return {"stats": stats, "logits": logits}


def fit(net, data, epochs, initial_epoch, validation_data, callbacks, verbose, steps_per_epoch,
        validation_steps):
    _check_compiled(net)
    hist = History()
    for cb in callbacks:
        cb.set_model(net)
        cb.on_train_begin()
    net.stop_training = False
    jlog = JsonlLogger(getattr(net, "log_jsonl", None), net.strategy.world.rank)
    for epoch in range(initial_epoch, epochs):
        t0 = time.time()
        for cb in callbacks:
            cb.on_epoch_begin(epoch)
        accum = _Accum(net)
        nb = 0
        seen = 0
        with nvtx_range(f"epoch {epoch}"):
            for x, y in _as_batches(data):
                n = x.shape[0]
                sl = net.strategy.local_slice(n)
                xs, ys = x[sl], y[sl]
                with nvtx_range("train_step"):
                    stats, logits = net.executor.train_batch(xs, ys, n)
                    gstats = net.strategy.reduce_and_apply(net.store, net.optimizer, 1.0, scalars=stats)
                accum.add(stats, logits, ys, gstats)
                nb += 1
                seen += n
                if steps_per_epoch is not None and nb >= steps_per_epoch:
                    break
        logs = accum.result()
        dt = max(time.time() - t0, 1e-9)
        if validation_data is not None:
            logs.update(_evaluate(net, validation_data, validation_steps, prefix="val_"))
        for cb in callbacks:
            cb.on_epoch_end(epoch, logs)          # (a callback may add to the logs, e.g. the learning rate)
        hist._append(epoch, logs)
        jlog.log("epoch", epoch=epoch, steps=nb, examples=seen, train_seconds=round(dt, 4),
                 examples_per_s=round(seen / dt, 1), **{k: float(v) for k, v in logs.items()})
        if verbose and net.strategy.world.rank == 0:
            body = " - ".join(f"{k}: {v:.4f}" for k, v in logs.items())
            print(f"Epoch {epoch + 1}/{epochs} - {time.time() - t0:.0f}s - {body}")
            sys.stdout.flush()
        if net.stop_training:
            break
    for cb in callbacks:
        cb.on_train_end()
    net.history = hist
    return hist


def _evaluate(net, data, steps, prefix="") -> dict:
    accum = _Accum(net)
    nb = 0
    for x, y in _as_batches(data):
        sl = net.strategy.local_slice(x.shape[0])
        xs, ys = x[sl], y[sl]
        stats, logits = net.executor.eval_batch(xs, ys)
        accum.add(stats, logits, ys)
        nb += 1
        if steps is not None and nb >= steps:
            break
    return accum.result(prefix)


def evaluate(net, data, steps, verbose, return_dict):
    _check_compiled(net)
    logs = _evaluate(net, data, steps)
    if verbose and net.strategy.world.rank == 0:
        print(" - ".join(f"{k}: {v:.4f}" for k, v in logs.items()))
    if return_dict:
        return logs
    return [logs["loss"]] + [logs[m.name] for m in net.metrics]


def predict(net, x):
    if isinstance(x, torch.Tensor):
        return net.executor.predict(x)
    outs = [net.executor.predict(b[0] if isinstance(b, tuple) else b) for b in x]
    return torch.cat(outs, 0)


def rebind_optimizer(net, optimizer) -> None:
    """Fresh optimizer (state) over the SAME trainable set, keeping the executor's plans and graphs —
    what a federated client needs at the start of every round."""
    opt = O.get(optimizer)
    if opt.store is not None:
        opt = opt.clone()
    net.optimizer = opt.bind(net.store, net.store.ranges(net.trainable_layer_indices()))
    net.store.grad.zero_()
    if hasattr(net.executor, "grad_hook"):
        net.executor.grad_hook = None
    after = getattr(net.strategy, "after_compile", None)
    if after is not None:
        after(net)              # an overlap schedule is tied to the optimizer object
