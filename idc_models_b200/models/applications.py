"""``tf.keras.applications``-style constructors and ``Sequential`` assembly.

    base = DenseNet121(input_shape=(50, 50, 3), include_top=False, weights=None)
    base.trainable = False
    model = Sequential([base, GlobalAveragePooling2D(), Dense(1)])

mirrors /root/reference/dist_model_tf_mobile.py:119-129 (and the dense / vgg / fed twins). ``weights``:
``None`` = Keras default random init (glorot-uniform); a path = Keras ``.h5`` weight file to import;
``'imagenet'`` needs a download, which is impossible offline, so it falls back to random init with a
warning.
"""
from __future__ import annotations

import warnings

from . import spec as S
from .network import Network


class BaseSpec:
    """Un-instantiated application (a spec + the requested weights); becomes ``model.base`` once placed
    in a :func:`Sequential`. ``trainable`` set here is carried over."""

    def __init__(self, arch: str, input_shape, weights=None):
        self.arch = arch
        self.spec = S.build_spec(arch, tuple(input_shape))
        self.weights = weights
        self.name = self.spec.name
        self._trainable = True
        self._bound: Network | None = None

    @property
    def layers(self):
        if self._bound is not None:
            return self._bound.base.layers
        return self.spec.layers

    @property
    def trainable(self):
        return self._bound.base.trainable if self._bound is not None else self._trainable

    @trainable.setter
    def trainable(self, v):
        self._trainable = bool(v)
        if self._bound is not None:
            self._bound.base.trainable = v


def _app(arch):
    def ctor(input_shape=(50, 50, 3), include_top=False, weights=None, **kw):
        if include_top:
            raise NotImplementedError("the reference only uses include_top=False")
        return BaseSpec(arch, input_shape, weights)
    ctor.__name__ = arch
    return ctor


DenseNet121 = _app("densenet121")
DenseNet169 = _app("densenet169")
DenseNet201 = _app("densenet201")
MobileNetV2 = _app("mobilenet_v2")
VGG16 = _app("vgg16")


# ---- single layers usable inside Sequential (secure_fed_model.py:84-94) ---------------------------
class _L:
    def __init__(self, kind, **attrs):
        self.kind, self.attrs = kind, attrs


def Conv2D(filters, kernel_size, strides=1, padding="valid", activation=None, use_bias=True, input_shape=None, **kw):
    k = kernel_size[0] if isinstance(kernel_size, (tuple, list)) else kernel_size
    s = strides[0] if isinstance(strides, (tuple, list)) else strides
    return _L("conv", filters=filters, k=k, stride=s, padding=padding, activation=activation,
              use_bias=use_bias, input_shape=input_shape)


def MaxPooling2D(pool_size=(2, 2), strides=None, **kw):
    k = pool_size[0] if isinstance(pool_size, (tuple, list)) else pool_size
    return _L("maxpool", k=k, stride=(strides or k))


def AveragePooling2D(pool_size=(2, 2), strides=None, **kw):
    k = pool_size[0] if isinstance(pool_size, (tuple, list)) else pool_size
    return _L("avgpool", k=k, stride=(strides or k))


def Dropout(rate, **kw):
    return _L("dropout", rate=rate)


def Flatten(**kw):
    return _L("flatten")


def GlobalAveragePooling2D(**kw):
    return _L("gap")


def Dense(units, activation=None, **kw):
    return _L("dense", units=units, activation=activation)


class _GrowingSequential:
    """``model = Sequential(); model.add(layer); ...`` (/root/reference/secure_fed_model.py:85-94): layers are
    collected until the model is first used (``compile``, ``summary``, ``get_weights``, ...); from then on the object
    is the built :class:`Network` in every respect (attribute access is forwarded to it)."""

    def __init__(self, **kw):
        object.__setattr__(self, "_items", [])
        object.__setattr__(self, "_kw", kw)
        object.__setattr__(self, "_net", None)

    def add(self, layer) -> None:
        if self._net is not None:
            raise RuntimeError("Sequential.add after the model has been built (compile / summary / weights were used)")
        self._items.append(layer)

    def build(self) -> Network:
        if self._net is None:
            if not self._items:
                raise ValueError("Sequential has no layers")
            object.__setattr__(self, "_net", Sequential(list(self._items), **self._kw))
        return self._net

    def __getattr__(self, name):
        return getattr(self.build(), name)

    def __setattr__(self, name, value):
        setattr(self.build(), name, value)


def Sequential(items=None, name: str = "sequential", device=None, seed: int = 0, executor: str = "auto",
               input_shape=None):
    """Chain an optional application base and single layers into one :class:`Network`. Without ``items`` the
    incremental Keras form is returned: ``model = Sequential(); model.add(Conv2D(...)); ...; model.compile(...)``."""
    if items is None:
        return _GrowingSequential(name=name, device=device, seed=seed, executor=executor, input_shape=input_shape)
    base: BaseSpec | None = None
    rest = list(items)
    if rest and isinstance(rest[0], BaseSpec):
        base = rest.pop(0)
        b = S._Builder.__new__(S._Builder)
        b.spec = S.ModelSpec(base.spec.name, list(base.spec.layers), base.spec.input_shape)
        b.last = base.spec.layers[-1].name
        ishape = base.spec.input_shape
    else:
        for it in rest:
            if input_shape is None and it.attrs.get("input_shape"):
                input_shape = it.attrs["input_shape"]
        if input_shape is None:
            raise ValueError("Sequential needs an input_shape (on the first layer or as an argument)")
        b = S._Builder(name, tuple(input_shape))
        ishape = tuple(input_shape)
    counts: dict[str, int] = {}

    def uniq(stem):
        i = counts.get(stem, 0)
        counts[stem] = i + 1
        return stem if i == 0 else f"{stem}_{i}"

    for it in rest:
        a = it.attrs
        if it.kind == "conv":
            b.conv(uniq("conv2d"), a["filters"], a["k"], a["stride"], a["padding"], a["use_bias"], a["activation"])
        elif it.kind == "maxpool":
            b.pool(uniq("max_pooling2d"), "maxpool", a["k"], a["stride"])
        elif it.kind == "avgpool":
            b.pool(uniq("average_pooling2d"), "avgpool", a["k"], a["stride"])
        elif it.kind == "dropout":
            b.dropout(uniq("dropout"), a["rate"])
        elif it.kind == "flatten":
            b.flatten(uniq("flatten"))
        elif it.kind == "gap":
            b.gap(uniq("global_average_pooling2d"))
        elif it.kind == "dense":
            b.dense(uniq("dense"), a["units"], a["activation"])
        else:
            raise ValueError(it.kind)
    nb = len(base.spec.layers) if base is not None else None
    net = Network(b.spec.layers, name, ishape, device, seed, executor, nb, base.name if base else None)
    if base is not None:
        net.base.trainable = base._trainable
        base._bound = net
        if base.weights == "imagenet":
            warnings.warn("weights='imagenet' needs a download; no network here -> random init. "
                          "Pass weights='<file>.h5' to import Keras weights.")
        elif isinstance(base.weights, str):
            from ..io import keras_h5
            keras_h5.load_into(net, base.weights, by_name=True, strict=False)
    return net


def build_transfer_model(arch: str, input_shape=(50, 50, 3), units: int = 1, weights=None, device=None,
                         seed: int = 0, executor: str = "auto") -> Network:
    """base(include_top=False) + GlobalAveragePooling2D + Dense(units), base frozen
    (dist_model_tf_mobile.py:119-129)."""
    base = BaseSpec(arch, input_shape, weights)
    base.trainable = False
    return Sequential([base, GlobalAveragePooling2D(), Dense(units)], name="sequential", device=device,
                      seed=seed, executor=executor)
