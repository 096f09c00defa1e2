"""Keras-shaped model object: ``Sequential([base, GlobalAveragePooling2D(), Dense(k)])`` with
``compile / fit / evaluate / predict / get_weights / set_weights / save_weights / load_weights`` and
per-layer ``trainable`` flags — the API surface the reference scripts drive
(/root/reference/dist_model_tf_mobile.py:117-160, fed_model.py:113-147, secure_fed_model.py:84-98,
131-154).

Execution is delegated to an executor chosen by device:
* CPU  -> :class:`models.reference.RefExecutor` (plain PyTorch fp32, autograd) — also the oracle;
* CUDA -> the fused sm_100a pipelines in :mod:`models.fused` (hand-written kernels, bf16 compute,
  fp32 master weights). There is no silent fallback on CUDA: a missing native library raises.
"""
from __future__ import annotations

import numpy as np
import torch

from ..nn.store import ParamStore
from ..parallel.strategy import current_strategy
from . import spec as S


class Layer:
    """One Keras layer: a name, a ``trainable`` flag and its weights."""

    def __init__(self, net: "Network", index: int, sp: S.LayerSpec):
        self._net, self.index, self.spec = net, index, sp
        self.name = sp.name
        self._trainable = True

    @property
    def trainable(self) -> bool:
        return self._trainable

    @trainable.setter
    def trainable(self, v: bool) -> None:
        self._trainable = bool(v)
        self._net._dirty = True

    @property
    def weights(self):
        return [self._net.store.view(p.name) for p in self.spec.params]

    def __repr__(self):
        return f"<Layer {self.index} {self.name} ({self.spec.kind}) trainable={self._trainable}>"


class LayerGroup:
    """A sub-model view (the ``base_model`` of the transfer scripts): ``.layers`` and a ``trainable``
    setter that flips every contained layer (dist_model_tf_mobile.py:122,141)."""

    def __init__(self, name: str, layers: list):
        self.name = name
        self.layers = layers

    @property
    def trainable(self) -> bool:
        return any(l.trainable for l in self.layers)

    @trainable.setter
    def trainable(self, v: bool) -> None:
        for l in self.layers:
            l.trainable = v


class Network:
    def __init__(self, layer_specs: list, name: str = "model", input_shape=(50, 50, 3), device=None,
                 seed: int = 0, executor: str = "auto", base_count: int | None = None,
                 base_name: str | None = None):
        self.name = name
        self.input_shape = tuple(input_shape)
        self.strategy = current_strategy()
        if device is None:
            device = self.strategy.world.device if self.strategy.name != "default" else (
                "cuda" if torch.cuda.is_available() else "cpu")
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.specs = layer_specs
        self.layers = [Layer(self, i, sp) for i, sp in enumerate(layer_specs)]
        self.store = ParamStore(layer_specs, self.device, seed)
        self.n_outputs = layer_specs[-1].out_shape[-1]
        nb = base_count if base_count is not None else len(self.layers)
        self.base = LayerGroup(base_name or name, self.layers[:nb])
        self._dirty = True
        self._executor_kind = executor
        self._executor = None
        self.optimizer = None
        self.loss = None
        self.metrics = []
        self.stop_training = False
        self.strategy.sync_variables(self.store)

    # ---- structure --------------------------------------------------------------------------
    def get_layer(self, name: str) -> Layer:
        for l in self.layers:
            if l.name == name:
                return l
        raise KeyError(name)

    def trainable_layer_indices(self) -> set:
        return {l.index for l in self.layers if l.trainable}

    @property
    def trainable_weights(self):
        """Keras order: per layer, its trainable variables (frozen layers and BN moving stats excluded)."""
        out = []
        for l in self.layers:
            if l.trainable:
                out += [self.store.view(p.name) for p in l.spec.params if p.trainable]
        return out

    @property
    def non_trainable_weights(self):
        out = []
        for l in self.layers:
            for p in l.spec.params:
                if not (l.trainable and p.trainable):
                    out.append(self.store.view(p.name))
        return out

    def count_params(self, trainable_only: bool = False) -> int:
        n = 0
        for l in self.layers:
            for p in l.spec.params:
                if trainable_only and not (l.trainable and p.trainable):
                    continue
                n += int(np.prod(p.shape))
        return n

    def summary(self, print_fn=print) -> None:
        print_fn(f'Model: "{self.name}"')
        for l in self.layers:
            n = sum(int(np.prod(p.shape)) for p in l.spec.params)
            print_fn(f"{l.index:4d} {l.name:36s} {l.spec.kind:8s} {str(l.spec.out_shape):18s} {n:10d}"
                     f"{'' if l.trainable else '  (frozen)'}")
        print_fn(f"Total params: {self.count_params():,}")
        print_fn(f"Trainable params: {self.count_params(True):,}")
        print_fn(f"Non-trainable params: {self.count_params() - self.count_params(True):,}")

    # ---- weights I/O (Keras layouts: Conv2D HWIO, Depthwise (kh,kw,C,1), Dense (in,out)) -----------
    @staticmethod
    def _to_keras(kind: str, t: torch.Tensor) -> np.ndarray:
        a = t.detach().to("cpu", torch.float32)
        if kind == "conv_kernel":
            a = a.permute(1, 2, 3, 0)
        elif kind == "dw_kernel":
            a = a.unsqueeze(-1)
        elif kind == "dense_kernel":
            a = a.t()
        return a.contiguous().numpy()

    @staticmethod
    def _from_keras(kind: str, a: np.ndarray) -> torch.Tensor:
        t = torch.as_tensor(np.asarray(a, dtype=np.float32))
        if kind == "conv_kernel":
            t = t.permute(3, 0, 1, 2)
        elif kind == "dw_kernel":
            t = t.reshape(t.shape[0], t.shape[1], t.shape[2])
        elif kind == "dense_kernel":
            t = t.t()
        return t.contiguous()

    def weight_names(self) -> list:
        return list(self.store.order)

    def get_weights(self) -> list:
        """List of numpy arrays in Keras ``model.get_weights()`` order and layouts."""
        return [self._to_keras(self.store.slots[n].kind, self.store.view(n)) for n in self.store.order]

    def set_weights(self, weights: list) -> None:
        if len(weights) != len(self.store.order):
            raise ValueError(f"expected {len(self.store.order)} arrays, got {len(weights)}")
        for n, a in zip(self.store.order, weights):
            s = self.store.slots[n]
            t = self._from_keras(s.kind, a)
            if tuple(t.shape) != tuple(s.shape):
                raise ValueError(f"{n}: shape {tuple(t.shape)} does not match {tuple(s.shape)}")
            self.store.view(n).copy_(t.to(self.device))
        self.store.refresh_shadow()

    def named_weights(self) -> dict:
        return {n: self._to_keras(self.store.slots[n].kind, self.store.view(n)) for n in self.store.order}

    def set_named_weights(self, d: dict, strict: bool = True) -> list:
        missing = []
        for n in self.store.order:
            if n not in d:
                missing.append(n)
                continue
            s = self.store.slots[n]
            t = self._from_keras(s.kind, d[n])
            if tuple(t.shape) != tuple(s.shape):
                raise ValueError(f"{n}: shape {tuple(t.shape)} does not match {tuple(s.shape)}")
            self.store.view(n).copy_(t.to(self.device))
        if strict and missing:
            raise KeyError(f"missing weights: {missing[:5]}{'...' if len(missing) > 5 else ''}")
        self.store.refresh_shadow()
        return missing

    def save_weights(self, path: str) -> None:
        from ..io import checkpoint
        self.strategy.aggregate_moving_stats(self.store)
        checkpoint.save_weights(self, path)

    def load_weights(self, path: str) -> None:
        from ..io import checkpoint
        checkpoint.load_weights(self, path)

    def save(self, path: str) -> None:
        """``model.save("m.h5")``: Keras full-model HDF5 (weights under /model_weights, model / training config,
        optimizer state). Any other extension writes the flat weights-only checkpoint."""
        self.strategy.aggregate_moving_stats(self.store)
        if path.endswith((".h5", ".hdf5")):
            from ..io import keras_h5
            keras_h5.save_model(self, path)
        else:
            from ..io import checkpoint
            checkpoint.save_weights(self, path)

    # ---- executor -------------------------------------------------------------------------------
    @property
    def executor(self):
        if self._executor is None:
            kind = self._executor_kind
            if kind == "auto":
                kind = "fused" if self.device.type == "cuda" else "reference"
            if kind in ("reference", "reference_bf16"):
                from .reference import RefExecutor
                self._executor = RefExecutor(self, emulate_bf16=(kind == "reference_bf16"))
            elif kind == "fused":
                from .fused import make_fused_executor
                self._executor = make_fused_executor(self)
            else:
                raise ValueError(kind)
        return self._executor

    # ---- Keras training API (implemented in train.engine) ---------------------------------------
    def compile(self, optimizer="rmsprop", loss=None, metrics=None, **kw) -> None:
        from ..train import engine
        engine.compile_model(self, optimizer, loss, metrics or [])

    def fit(self, x, epochs: int = 1, initial_epoch: int = 0, validation_data=None, callbacks=None,
            verbose: int = 1, steps_per_epoch: int | None = None, validation_steps: int | None = None):
        from ..train import engine
        return engine.fit(self, x, epochs, initial_epoch, validation_data, callbacks or [], verbose,
                          steps_per_epoch, validation_steps)

    def evaluate(self, x, steps: int | None = None, verbose: int = 0, return_dict: bool = False):
        from ..train import engine
        return engine.evaluate(self, x, steps, verbose, return_dict)

    def predict(self, x) -> torch.Tensor:
        from ..train import engine
        return engine.predict(self, x)

    def train_on_batch(self, x, y) -> dict:
        from ..train import engine
        return engine.train_on_batch(self, x, y)

    def to(self, device) -> "Network":
        device = torch.device(device)
        if device != self.device:
            self.device = device
            self.store.to(device)
            self._executor = None
            self._dirty = True
        return self


def clone_model(net: Network, device=None, seed: int = 1) -> Network:
    """``tf.keras.models.clone_model``: same architecture and trainable flags, FRESH weights
    (fed_model.py:200)."""
    c = Network(net.specs, net.name, net.input_shape, device or net.device, seed, net._executor_kind,
                len(net.base.layers), net.base.name)
    for a, b in zip(c.layers, net.layers):
        a.trainable = b.trainable
    return c
