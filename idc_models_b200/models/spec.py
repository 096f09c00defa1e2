"""Architecture descriptions in Keras layer order.

The reference builds its networks with ``tf.keras.applications`` and then freezes
``base_model.layers[:fine_tune_at]`` (/root/reference/dist_model_tf_dense.py:131,158-163,
dist_model_tf_mobile.py:119,146-151, dist_model_tf_vgg.py:119,146-151, fed_model.py:113,145-146).
``fine_tune_at`` indexes Keras' layer enumeration, so this module restates each application as a
flat list of :class:`LayerSpec` in exactly that order (index 0 = InputLayer), with Keras layer and
weight names (needed for ``.h5`` import/export) and a tiny dataflow IR (each layer names its inputs)
that both executors walk: the PyTorch fp32 reference interpreter and the fused sm_100a pipelines.

Layer counts (no top): DenseNet121 427, DenseNet169 595, DenseNet201 707, MobileNetV2 155, VGG16 19.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any


@dataclass
class ParamSpec:
    name: str            # Keras weight name without the ':0', e.g. 'conv1/conv/kernel'
    shape: tuple         # OUR storage shape
    kind: str            # conv_kernel | dw_kernel | dense_kernel | bias | gamma | beta | moving_mean | moving_var
    trainable: bool = True   # False for BN moving statistics
    fan_in: int = 0
    fan_out: int = 0


@dataclass
class LayerSpec:
    name: str
    kind: str                      # input|zeropad|conv|dwconv|bn|act|maxpool|avgpool|concat|add|gap|dense|dropout|flatten
    inputs: list = field(default_factory=list)
    attrs: dict = field(default_factory=dict)
    params: list = field(default_factory=list)   # list[ParamSpec]
    out_shape: tuple = ()          # (H, W, C) or (C,)


@dataclass
class ModelSpec:
    name: str                      # Keras model name: densenet121 | mobilenetv2_1.00_224 | vgg16 | ...
    layers: list                   # list[LayerSpec]
    input_shape: tuple

    def index(self, layer_name: str) -> int:
        for i, l in enumerate(self.layers):
            if l.name == layer_name:
                return i
        raise KeyError(layer_name)

    @property
    def output(self) -> LayerSpec:
        return self.layers[-1]

    def param_count(self, trainable_only: bool = False) -> int:
        n = 0
        for l in self.layers:
            for p in l.params:
                if trainable_only and not p.trainable:
                    continue
                k = 1
                for s in p.shape:
                    k *= s
                n += k
        return n


class _Builder:
    def __init__(self, name: str, input_shape: tuple):
        self.spec = ModelSpec(name, [], input_shape)
        self.last = None
        self._add(LayerSpec("input_1", "input", [], {}, [], tuple(input_shape)))

    def _add(self, l: LayerSpec) -> str:
        self.spec.layers.append(l)
        self.last = l.name
        return l.name

    def shape(self, name: str) -> tuple:
        return self.spec.layers[self.spec.index(name)].out_shape

    def zeropad(self, name, pad, src=None):
        src = src or self.last
        h, w, c = self.shape(src)
        (pt, pb), (pl, pr) = pad
        return self._add(LayerSpec(name, "zeropad", [src], {"pad": pad}, [], (h + pt + pb, w + pl + pr, c)))

    def conv(self, name, cout, k, stride=1, padding="same", use_bias=False, activation=None, src=None,
             weight_prefix=None):
        src = src or self.last
        h, w, c = self.shape(src)
        if padding == "same":
            ho, wo = -(-h // stride), -(-w // stride)
        else:
            ho, wo = (h - k) // stride + 1, (w - k) // stride + 1
        wp = weight_prefix or name
        ps = [ParamSpec(f"{wp}/kernel", (cout, k, k, c), "conv_kernel", True, k * k * c, k * k * cout)]
        if use_bias:
            ps.append(ParamSpec(f"{wp}/bias", (cout,), "bias"))
        return self._add(LayerSpec(name, "conv", [src], {"k": k, "stride": stride, "padding": padding,
                                                        "use_bias": use_bias, "activation": activation},
                                   ps, (ho, wo, cout)))

    def dwconv(self, name, k, stride=1, padding="same", src=None):
        src = src or self.last
        h, w, c = self.shape(src)
        if padding == "same":
            ho, wo = -(-h // stride), -(-w // stride)
        else:
            ho, wo = (h - k) // stride + 1, (w - k) // stride + 1
        ps = [ParamSpec(f"{name}/depthwise_kernel", (k, k, c), "dw_kernel", True, k * k, k * k)]
        return self._add(LayerSpec(name, "dwconv", [src], {"k": k, "stride": stride, "padding": padding},
                                   ps, (ho, wo, c)))

    def bn(self, name, eps, momentum=0.99, src=None):
        src = src or self.last
        shp = self.shape(src)
        c = shp[-1]
        ps = [ParamSpec(f"{name}/gamma", (c,), "gamma"), ParamSpec(f"{name}/beta", (c,), "beta"),
              ParamSpec(f"{name}/moving_mean", (c,), "moving_mean", False),
              ParamSpec(f"{name}/moving_variance", (c,), "moving_var", False)]
        return self._add(LayerSpec(name, "bn", [src], {"eps": eps, "momentum": momentum}, ps, shp))

    def act(self, name, max_value=None, src=None):
        src = src or self.last
        return self._add(LayerSpec(name, "act", [src], {"max_value": max_value}, [], self.shape(src)))

    def pool(self, name, kind, k, stride, src=None):
        src = src or self.last
        h, w, c = self.shape(src)
        return self._add(LayerSpec(name, kind, [src], {"k": k, "stride": stride}, [],
                                   ((h - k) // stride + 1, (w - k) // stride + 1, c)))

    def concat(self, name, srcs):
        h, w, _ = self.shape(srcs[0])
        c = sum(self.shape(s)[-1] for s in srcs)
        return self._add(LayerSpec(name, "concat", list(srcs), {}, [], (h, w, c)))

    def add(self, name, srcs):
        return self._add(LayerSpec(name, "add", list(srcs), {}, [], self.shape(srcs[0])))

    def gap(self, name, src=None):
        src = src or self.last
        return self._add(LayerSpec(name, "gap", [src], {}, [], (self.shape(src)[-1],)))

    def flatten(self, name, src=None):
        src = src or self.last
        n = 1
        for s in self.shape(src):
            n *= s
        return self._add(LayerSpec(name, "flatten", [src], {}, [], (n,)))

    def dropout(self, name, rate, src=None):
        src = src or self.last
        return self._add(LayerSpec(name, "dropout", [src], {"rate": rate}, [], self.shape(src)))

    def dense(self, name, units, activation=None, src=None):
        src = src or self.last
        (cin,) = self.shape(src)
        ps = [ParamSpec(f"{name}/kernel", (units, cin), "dense_kernel", True, cin, units),
              ParamSpec(f"{name}/bias", (units,), "bias")]
        return self._add(LayerSpec(name, "dense", [src], {"units": units, "activation": activation}, ps,
                                   (units,)))


# ---------------------------------------------------------------------------------------------
# DenseNet (keras.applications.densenet) — dist_model_tf_dense.py:131
# ---------------------------------------------------------------------------------------------
DENSENET_BLOCKS = {"densenet121": (6, 12, 24, 16), "densenet169": (6, 12, 32, 32),
                   "densenet201": (6, 12, 48, 32),
                   # not a Keras application: a shallow variant used by the numerics tests, where the
                   # fused bf16 pipeline can be compared tightly against the fp32 oracle
                   "densenet_tiny": (2, 2, 2, 2)}
DENSENET_EPS = 1.001e-5
GROWTH = 32


def densenet(arch: str = "densenet121", input_shape=(50, 50, 3)) -> ModelSpec:
    blocks = DENSENET_BLOCKS[arch]
    b = _Builder(arch, input_shape)
    b.zeropad("zero_padding2d", ((3, 3), (3, 3)))
    b.conv("conv1/conv", 64, 7, 2, "valid")
    b.bn("conv1/bn", DENSENET_EPS)
    b.act("conv1/relu")
    b.zeropad("zero_padding2d_1", ((1, 1), (1, 1)))
    b.pool("pool1", "maxpool", 3, 2)
    x = b.last
    for bi, nlayers in enumerate(blocks):
        stage = f"conv{bi + 2}"
        for li in range(nlayers):
            n = f"{stage}_block{li + 1}"
            b.bn(f"{n}_0_bn", DENSENET_EPS, src=x)
            b.act(f"{n}_0_relu")
            b.conv(f"{n}_1_conv", 4 * GROWTH, 1)
            b.bn(f"{n}_1_bn", DENSENET_EPS)
            b.act(f"{n}_1_relu")
            y = b.conv(f"{n}_2_conv", GROWTH, 3, 1, "same")
            x = b.concat(f"{n}_concat", [x, y])
        if bi != len(blocks) - 1:
            t = f"pool{bi + 2}"
            c = b.shape(x)[-1]
            b.bn(f"{t}_bn", DENSENET_EPS, src=x)
            b.act(f"{t}_relu")
            b.conv(f"{t}_conv", c // 2, 1)
            x = b.pool(f"{t}_pool", "avgpool", 2, 2)
    b.bn("bn", DENSENET_EPS, src=x)
    b.act("relu")
    return b.spec


# ---------------------------------------------------------------------------------------------
# MobileNetV2 alpha=1.0 (keras.applications.mobilenet_v2, TF 2.1 layer order) — dist_model_tf_mobile.py:119
# ---------------------------------------------------------------------------------------------
MBV2_EPS = 1e-3
MBV2_MOM = 0.999
MBV2_CFG = [  # t, c, n, s
    (1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]


def correct_pad(size: int, k: int = 3) -> tuple[int, int]:
    """keras_applications.correct_pad for one spatial dim: even input -> (k//2 - 1, k//2), odd -> (k//2, k//2)."""
    adjust = 1 - size % 2
    c = k // 2
    return c - adjust, c


MBV2_CFGS = {"mobilenetv2_1.00_224": MBV2_CFG,
             # shallow test variant (see densenet_tiny)
             "mobilenetv2_tiny": [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 2, 2)]}


def mobilenet_v2(input_shape=(50, 50, 3), name: str = "mobilenetv2_1.00_224") -> ModelSpec:
    b = _Builder(name, input_shape)
    MBV2_CFG = MBV2_CFGS[name]  # noqa: N806 - shadows the module constant on purpose
    h, w, _ = input_shape
    b.zeropad("Conv1_pad", (correct_pad(h), correct_pad(w)))
    b.conv("Conv1", 32, 3, 2, "valid")
    b.bn("bn_Conv1", MBV2_EPS, MBV2_MOM)
    b.act("Conv1_relu", 6.0)
    x = b.last
    block_id = 0
    for t, c, n, s in MBV2_CFG:
        for i in range(n):
            stride = s if i == 0 else 1
            prefix = "expanded_conv_" if block_id == 0 else f"block_{block_id}_"
            inp = x
            cin = b.shape(inp)[-1]
            if block_id != 0:
                b.conv(prefix + "expand", t * cin, 1, src=inp)
                b.bn(prefix + "expand_BN", MBV2_EPS, MBV2_MOM)
                b.act(prefix + "expand_relu", 6.0)
            if stride == 2:
                hh, ww, _ = b.shape(b.last)
                b.zeropad(prefix + "pad", (correct_pad(hh), correct_pad(ww)))
            b.dwconv(prefix + "depthwise", 3, stride, "same" if stride == 1 else "valid",
                     src=b.last if (block_id != 0 or stride == 2) else inp)
            b.bn(prefix + "depthwise_BN", MBV2_EPS, MBV2_MOM)
            b.act(prefix + "depthwise_relu", 6.0)
            b.conv(prefix + "project", c, 1)
            x = b.bn(prefix + "project_BN", MBV2_EPS, MBV2_MOM)
            if cin == c and stride == 1:
                x = b.add(prefix + "add", [inp, x])
            block_id += 1
    b.conv("Conv_1", 1280, 1, src=x)
    b.bn("Conv_1_bn", MBV2_EPS, MBV2_MOM)
    b.act("out_relu", 6.0)
    return b.spec


# ---------------------------------------------------------------------------------------------
# VGG16 (keras.applications.vgg16) — dist_model_tf_vgg.py:119, fed_model.py:113
# ---------------------------------------------------------------------------------------------
def vgg16(input_shape=(50, 50, 3)) -> ModelSpec:
    b = _Builder("vgg16", input_shape)
    for bi, (n, c) in enumerate(((2, 64), (2, 128), (3, 256), (3, 512), (3, 512))):
        for i in range(n):
            b.conv(f"block{bi + 1}_conv{i + 1}", c, 3, 1, "same", use_bias=True, activation="relu")
        b.pool(f"block{bi + 1}_pool", "maxpool", 2, 2)
    return b.spec


# ---------------------------------------------------------------------------------------------
# The 10x10 CNN of the secure-FL script — secure_fed_model.py:84-98 (1,937 parameters, 6 tensors)
# ---------------------------------------------------------------------------------------------
def small_cnn(input_shape=(10, 10, 3)) -> ModelSpec:
    b = _Builder("sequential", input_shape)
    b.conv("conv2d", 32, 3, 2, "valid", use_bias=True, activation="relu")
    b.pool("max_pooling2d", "maxpool", 2, 2)
    b.dropout("dropout", 0.25)
    b.flatten("flatten")
    b.dense("dense", 8, activation="relu")
    b.dropout("dropout_1", 0.5)
    b.dense("dense_1", 1)
    return b.spec


ARCHS = {
    "densenet121": lambda shape=(50, 50, 3): densenet("densenet121", shape),
    "densenet169": lambda shape=(50, 50, 3): densenet("densenet169", shape),
    "densenet201": lambda shape=(50, 50, 3): densenet("densenet201", shape),
    "mobilenet_v2": lambda shape=(50, 50, 3): mobilenet_v2(shape),
    "densenet_tiny": lambda shape=(50, 50, 3): densenet("densenet_tiny", shape),
    "mobilenet_v2_tiny": lambda shape=(50, 50, 3): mobilenet_v2(shape, "mobilenetv2_tiny"),
    "vgg16": lambda shape=(50, 50, 3): vgg16(shape),
    "small_cnn": lambda shape=(10, 10, 3): small_cnn(shape),
}


def build_spec(arch: str, input_shape=None) -> ModelSpec:
    if arch not in ARCHS:
        raise ValueError(f"unknown architecture {arch!r}; choose from {sorted(ARCHS)}")
    return ARCHS[arch](input_shape) if input_shape is not None else ARCHS[arch]()


def head_spec(base: ModelSpec, units: int) -> list:
    """GlobalAveragePooling2D + Dense(units) appended by every transfer script
    (dist_model_tf_mobile.py:123-129). Returned as extra layers after the base."""
    b = _Builder.__new__(_Builder)
    b.spec = ModelSpec(base.name, list(base.layers), base.input_shape)
    b.last = base.layers[-1].name
    n0 = len(base.layers)
    b.gap("global_average_pooling2d")
    b.dense("dense", units)
    return b.spec.layers[n0:]
