"""Keras ``.h5`` weight files without h5py/libhdf5 (neither exists on this image): import and export through
the minimal HDF5 subset in ``libidcb200_host.so`` (csrc/host/hdf5.cpp). SURVEY.md 2.B B4, 5.4, 8.16; the
reference obtains its ImageNet weights through ``tf.keras.applications(..., weights='imagenet')``
(/root/reference/dist_model_tf_mobile.py:119), i.e. from files of exactly this layout.

Layout written (``model.save_weights('x.h5')`` of ``Sequential([base, GAP, Dense])``):
  root attrs  layer_names = [<base name>, 'global_average_pooling2d', 'dense'], backend, keras_version
  /<layer>    attr weight_names = ['conv1/conv/kernel:0', ...]; datasets /<layer>/<weight name>
Tensors use Keras layouts: Conv2D HWIO, DepthwiseConv2D (kh, kw, C, 1), Dense (in, out), BN gamma/beta/
moving_mean/moving_variance; float32 little-endian, contiguous.

Reading accepts both this nested form and the flat per-layer form of the published
``*_weights_tf_dim_ordering_tf_kernels_notop.h5`` files: weights are matched by the suffix
``<weight name>:0`` of the dataset path.
"""
from __future__ import annotations

import ctypes as C
import json

import numpy as np

from ..ops import _lib


def _lib_h():
    lib = _lib.host_lib()
    lib.idcbh_h5w_open.restype = C.c_void_p
    lib.idcbh_h5r_open.restype = C.c_void_p
    lib.idcbh_h5r_list.restype = C.c_longlong
    return lib


def _err(lib) -> str:
    return lib.idcbh_last_error().decode(errors="replace")


def write_h5(path: str, datasets: dict, attrs: list) -> None:
    """datasets: {hdf5 path: float32 ndarray}; attrs: [(group path, name, [strings], scalar)]."""
    lib = _lib_h()
    h = C.c_void_p(lib.idcbh_h5w_open(path.encode()))
    keep = []
    for p, a in datasets.items():
        a = np.ascontiguousarray(a, dtype=np.float32)
        keep.append(a)
        dims = (C.c_longlong * max(a.ndim, 1))(*(a.shape if a.ndim else (1,)))
        if lib.idcbh_h5w_dataset_f32(h, p.encode(), a.ctypes.data_as(C.c_void_p), dims, C.c_int(a.ndim)) != 0:
            raise IOError(_err(lib))
    for group, name, values, scalar in attrs:
        arr = (C.c_char_p * max(len(values), 1))(*[v.encode() for v in values])
        if lib.idcbh_h5w_attr_strs(h, group.encode(), name.encode(), arr, C.c_int(len(values)), C.c_int(1 if scalar else 0)) != 0:
            raise IOError(_err(lib))
    if lib.idcbh_h5w_close(h) != 0:
        raise IOError(_err(lib))


def read_h5(path: str):
    """-> (datasets {path: ndarray}, attrs {(object path, name): [values]})"""
    lib = _lib_h()
    h = lib.idcbh_h5r_open(path.encode())
    if not h:
        raise IOError(_err(lib))
    h = C.c_void_p(h)
    try:
        need = lib.idcbh_h5r_list(h, None, C.c_longlong(0))
        buf = C.create_string_buffer(int(need))
        lib.idcbh_h5r_list(h, buf, C.c_longlong(need))
        meta = json.loads(buf.value.decode())
        data = {}
        for d in meta["datasets"]:
            if not d["readable"]:
                raise IOError(f"{path}: dataset {d['path']}: {d['why']}")
            if d["class"] != 1:
                continue
            a = np.empty(d["dims"] if d["dims"] else (), dtype=np.float32)
            if lib.idcbh_h5r_read_f32(h, d["path"].encode(), a.ctypes.data_as(C.c_void_p), C.c_longlong(a.size)) != 0:
                raise IOError(_err(lib))
            data[d["path"]] = a
        attrs = {(a["object"], a["name"]): a["values"] for a in meta["attrs"]}
        return data, attrs
    finally:
        lib.idcbh_h5r_close(h)


def _keras_layer_names(net) -> list:
    """top-level layer names of Sequential([base, ...head layers])"""
    nb = len(net.base.layers)
    if nb == len(net.layers):       # no nested base: every layer is top level
        return [l.name for l in net.layers if l.spec.kind != "input"]
    return [net.base.name] + [l.name for l in net.layers[nb:]]


def save_weights(net, path: str) -> None:
    if net.strategy.world.rank != 0:
        return
    named = net.named_weights()
    nb = len(net.base.layers)
    nested = nb != len(net.layers)
    layer_names = _keras_layer_names(net)
    groups: dict[str, list] = {n: [] for n in layer_names}
    for l in net.layers:
        if l.spec.kind == "input":
            continue
        top = net.base.name if (nested and l.index < nb) else l.name
        for p in l.spec.params:
            groups[top].append(p.name + ":0")
    datasets = {}
    for top, names in groups.items():
        for wn in names:
            datasets[f"{top}/{wn}"] = named[wn[:-2]]
    attrs = [("", "backend", ["tensorflow"], True), ("", "keras_version", ["2.2.4-tf"], True)]
    # Keras splits attributes above ~64 KB into name0, name1, ...
    def add(group, name, values):
        size = max([len(v) for v in values] + [1]) * max(len(values), 1)
        if size <= 60000:
            attrs.append((group, name, values, False))
            return
        per = max(1, 60000 // max(len(v) for v in values))
        for i in range(0, len(values), per):
            attrs.append((group, f"{name}{i // per}", values[i:i + per], False))
    add("", "layer_names", layer_names)
    for top, names in groups.items():
        add(top, "weight_names", names)
    write_h5(path, datasets, attrs)


def load_into(net, path: str, by_name: bool = True, strict: bool = True) -> list:
    """Load every weight whose Keras name appears in the file; returns the names that were missing."""
    data, _ = read_h5(path)
    # A Keras file stores weight <w> of top-level layer <top> at /<top>/<w>:0 (under /model_weights for a
    # model.save file). The same short name can also end a longer path ("bn/gamma" ends ".../conv1/bn/gamma"),
    # so the match with the fewest leading groups wins and a tie is an error rather than a guess that depends
    # on the order the file lists its groups in.
    paths = {p[:-2].strip("/"): a for p, a in data.items() if p.endswith(":0")}
    by_tail: dict[str, list] = {}
    for p in paths:
        by_tail.setdefault(p.rsplit("/", 1)[-1], []).append(p)
    found = {}
    for n in net.weight_names():
        cands = [p for p in by_tail.get(n.rsplit("/", 1)[-1], ()) if p == n or p.endswith("/" + n)]
        if not cands:
            continue
        depth = lambda p: p.count("/") - n.count("/")      # noqa: E731 - number of leading groups
        best = min(depth(p) for p in cands)
        top = [p for p in cands if depth(p) == best]
        if len(top) > 1:
            raise KeyError(f"{path}: weight name {n!r} is ambiguous: {top[:3]}")
        found[n] = paths[top[0]]
    missing = net.set_named_weights(found, strict=False)
    if strict and missing:
        raise KeyError(f"{path}: missing weights {missing[:5]}{'...' if len(missing) > 5 else ''}")
    return missing


# ---------------------------------------------------------------------------------------------------------
# Full-model file (``model.save("m.h5")``): /model_weights/... + model_config / training_config attributes +
# /optimizer_weights. ``load_into`` already accepts such a file (weights are matched by the
# tail of their path); ``load_optimizer_state`` restores the optimizer slots.
# ---------------------------------------------------------------------------------------------------------
def _optimizer_entries(net) -> tuple:
    """(names, {name: array}) in Keras' naming: ``<Opt>/iter:0``, hyper-parameters, then one slot per trainable
    variable (``<Opt>/<var>/rms:0`` for RMSprop, ``/momentum:0`` for SGD, ``/m:0`` and ``/v:0`` for Adam)."""
    import numpy as np
    opt = net.optimizer
    if opt is None:
        return [], {}
    cls = {"rmsprop": "RMSprop", "sgd": "SGD", "adam": "Adam"}[opt.kind]
    names, data = [], {}

    def put(n, a):
        names.append(n)
        data[n] = np.asarray(a, dtype=np.float32)

    put(f"{cls}/iter:0", [float(opt.iterations)])
    put(f"{cls}/learning_rate:0", [float(opt.learning_rate)])
    slots = {"rmsprop": [("rms", "s1")], "sgd": [("momentum", "s1")], "adam": [("m", "s1"), ("v", "s2")]}[opt.kind]
    st = net.store
    ranges = opt.ranges
    for wn in st.order:
        sl = st.slots[wn]
        if not sl.trainable_kind or not any(b <= sl.offset < e for b, e in ranges):
            continue
        for suffix, attr in slots:
            buf = getattr(opt, attr)
            if buf is None:
                continue
            put(f"{cls}/{wn}/{suffix}:0", net._to_keras(sl.kind, st._logical(buf, sl)))
    return names, data


def save_model(net, path: str) -> None:
    """Keras full-model HDF5 layout: the weights-only tree under ``/model_weights``, JSON ``model_config`` and
    ``training_config`` root attributes, optimizer state under ``/optimizer_weights``."""
    import json
    if net.strategy.world.rank != 0:
        return
    named = net.named_weights()
    nb = len(net.base.layers)
    nested = nb != len(net.layers)
    layer_names = _keras_layer_names(net)
    groups: dict = {n: [] for n in layer_names}
    for l in net.layers:
        if l.spec.kind == "input":
            continue
        top = net.base.name if (nested and l.index < nb) else l.name
        for p in l.spec.params:
            groups[top].append(p.name + ":0")
    datasets = {}
    for top, names in groups.items():
        for wn in names:
            datasets[f"model_weights/{top}/{wn}"] = named[wn[:-2]]
    config = {"class_name": "Sequential",
              "config": {"name": net.name, "layers": [{"class_name": l.spec.kind, "config": {"name": l.name, "trainable": bool(l.trainable)}}
                                                       for l in net.layers]},
              "input_shape": list(net.input_shape), "base": net.base.name}
    opt = net.optimizer
    training = {"loss": getattr(net.loss, "kind", None), "metrics": [getattr(m, "name", str(m)) for m in (net.metrics or [])],
                "optimizer_config": ({"class_name": opt.kind, "config": {"learning_rate": float(opt.learning_rate)}} if opt else None)}
    attrs = [("", "backend", ["tensorflow"], True), ("", "keras_version", ["2.2.4-tf"], True),
             ("", "model_config", [json.dumps(config)], True), ("", "training_config", [json.dumps(training)], True),
             ("model_weights", "backend", ["tensorflow"], True), ("model_weights", "keras_version", ["2.2.4-tf"], True)]

    def add(group, name, values):
        size = max([len(v) for v in values] + [1]) * max(len(values), 1)
        if size <= 60000:
            attrs.append((group, name, values, False))
            return
        per = max(1, 60000 // max(len(v) for v in values))
        for i in range(0, len(values), per):
            attrs.append((group, f"{name}{i // per}", values[i:i + per], False))
    add("model_weights", "layer_names", layer_names)
    for top, names in groups.items():
        add(f"model_weights/{top}", "weight_names", names)
    onames, odata = _optimizer_entries(net)
    if onames:
        for n in onames:
            datasets[f"optimizer_weights/{n}"] = odata[n]
        add("optimizer_weights", "weight_names", onames)
    write_h5(path, datasets, attrs)


def load_optimizer_state(net, path: str) -> int:
    """Restore the optimizer slots written by :func:`save_model` into the compiled model's optimizer (same optimizer
    class). Returns the number of slot tensors restored."""
    data, _ = read_h5(path)
    opt = net.optimizer
    if opt is None:
        raise RuntimeError("compile the model before loading optimizer state")
    cls = {"rmsprop": "RMSprop", "sgd": "SGD", "adam": "Adam"}[opt.kind]
    slots = {"rmsprop": [("rms", "s1")], "sgd": [("momentum", "s1")], "adam": [("m", "s1"), ("v", "s2")]}[opt.kind]
    st = net.store
    n = 0
    it = data.get(f"optimizer_weights/{cls}/iter:0")
    if it is not None:
        opt.iterations = int(round(float(it.reshape(-1)[0])))
    for wn in st.order:
        sl = st.slots[wn]
        if not sl.trainable_kind:
            continue
        for suffix, attr in slots:
            a = data.get(f"optimizer_weights/{cls}/{wn}/{suffix}:0")
            buf = getattr(opt, attr)
            if a is None or buf is None:
                continue
            st._logical(buf, sl).copy_(net._from_keras(sl.kind, a).to(buf.device))
            n += 1
    return n
