"""Name-for-name front door for code written against the reference's imports.

The five upstream scripts use a small, fixed part of TensorFlow / Keras, TensorFlow Federated, ``phe`` and
scikit-learn. This module lays this framework's implementations out under the same dotted names, so a
script written the way the reference's are ports by changing its import lines::

    # was:  import tensorflow as tf; import tensorflow_federated as tff; from phe import paillier
    from idc_models_b200.compat import tf, tff, tfds, paillier, roc_auc_score

    strategy = tf.distribute.MirroredStrategy()
    with strategy.scope():
        base_model = tf.keras.applications.MobileNetV2(input_shape=(50, 50, 3), include_top=False, weights=None)
        base_model.trainable = False
        model = tf.keras.Sequential([base_model, tf.keras.layers.GlobalAveragePooling2D(), tf.keras.layers.Dense(1)])
        model.compile(optimizer=tf.keras.optimizers.RMSprop(lr=1e-4),
                      loss=tf.keras.losses.BinaryCrossentropy(from_logits=True), metrics=['accuracy'])

Only what the reference touches is here (/root/reference/dist_model_tf_mobile.py:1-20,105-170, fed_model.py:1-30,
186-236, secure_fed_model.py:1-20,85-121); it is a map of names, not an emulation of TensorFlow. Tensors are
``torch.Tensor`` / ``numpy.ndarray``, datasets are :class:`idc_models_b200.data.dataset.Dataset`.
"""
from __future__ import annotations

import types as _types

from .data import dataset as _dataset
from .fed import client_data as _client_data
from .fed import fedavg as _fedavg
from .fed import paillier as _paillier
from .models import applications as _apps
from .models import network as _network
from .optim import optimizers as _optim
from .parallel import strategy as _strategy
from .train import engine as _engine
from .train import losses as _losses
from .train import metrics as _metrics


def _ns(**kw) -> _types.SimpleNamespace:
    return _types.SimpleNamespace(**kw)


def _read_file(path: str) -> bytes:
    """tf.io.read_file"""
    with open(path, "rb") as f:
        return f.read()


def _decode_png(data: bytes, channels: int = 3):
    """tf.image.decode_png(channels=3) followed by convert_image_dtype(float32): float32 [H, W, 3] in [0, 1] at the
    image's own size (the two reference calls are one step here; ``convert_image_dtype`` is then the identity)."""
    import numpy as np
    from PIL import Image
    import io
    if channels != 3:
        raise ValueError("only channels=3 is supported (what the reference uses)")
    lib = _dataset._native_png()
    if lib and len(data) >= 24 and data[:8] == b"\x89PNG\r\n\x1a\n" and data[12:16] == b"IHDR":
        w, h = int.from_bytes(data[16:20], "big"), int.from_bytes(data[20:24], "big")   # the image's own size
        if 0 < w <= 16384 and 0 < h <= 16384:
            out = np.empty((h, w, 3), np.float32)
            if lib.idcbh_png_decode_f32(bytes(data), len(data), out.ctypes.data, h, w, None, None) == 0:
                return out
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"), dtype=np.float32) / 255.0


def _convert_image_dtype(img, dtype=None):
    import numpy as np
    a = np.asarray(img)
    return a.astype(np.float32) / 255.0 if a.dtype == np.uint8 else a.astype(np.float32)


def _resize(img, size):
    """tf.image.resize (bilinear, half-pixel centres, no antialias) of one [H, W, C] float image"""
    import numpy as np
    import torch
    a = np.asarray(img, dtype=np.float32)
    if a.shape[0] == size[0] and a.shape[1] == size[1]:
        return a
    t = torch.from_numpy(a).permute(2, 0, 1).unsqueeze(0)
    t = torch.nn.functional.interpolate(t, size=tuple(size), mode="bilinear", align_corners=False)
    return t.squeeze(0).permute(1, 2, 0).contiguous().numpy()


def _strings_split(s, sep=None):
    """tf.strings.split of ONE string: the list of parts (``parts[-2] == '1'`` is the reference's label rule)"""
    return str(s).split(sep)


def _cast(x, dtype):
    import numpy as np
    import torch
    if isinstance(x, torch.Tensor):
        return x.to(torch.float64 if dtype in (np.float64, float) else torch.float32 if dtype is np.float32 else dtype)
    return np.asarray(x).astype(dtype)


def _map_structure(fn, *structures):
    """tf.nest.map_structure over nested lists / tuples / dicts"""
    first = structures[0]
    if isinstance(first, dict):
        return {k: _map_structure(fn, *[s[k] for s in structures]) for k in first}
    if isinstance(first, (list, tuple)):
        out = [_map_structure(fn, *parts) for parts in zip(*structures)]
        return type(first)(out) if not hasattr(first, "_fields") else type(first)(*out)
    return fn(*structures)


def _py_func(func, inp, Tout=None, **kw):
    """tf.compat.v1.py_func: there is no graph here, so it is the call itself"""
    return func(*inp)


def _tfds_load(name="cifar10", split=None, data_dir=None, with_info=False, as_supervised=True, n_synthetic=None, seed=0):
    """tfds.load('cifar10', with_info=True, as_supervised=True) of /root/reference/dist_model_tf_dense.py:120: an
    on-disk copy under ``data_dir`` (``cifar-10-batches-bin`` / ``-py``, searched like the dense script does) or,
    without one, the synthetic stand-in of the same shape (there is nothing to download from here). uint8 images."""
    if name != "cifar10":
        raise ValueError("only 'cifar10' (the dataset the reference loads) is available")
    import os
    from .data import cifar, synthetic
    if n_synthetic is None:
        n_synthetic = int(os.environ.get("IDCB_TFDS_SYNTHETIC_N", "50000"))
    data_dir = data_dir or os.environ.get("TFDS_DATA_DIR")
    found = cifar.find_cifar10(data_dir) if data_dir else None
    if found is not None:
        (xtr, ytr), (xte, yte) = cifar.load_cifar10(found)
        ds = {"train": _dataset.Dataset.from_tensor_slices((xtr, ytr)), "test": _dataset.Dataset.from_tensor_slices((xte, yte))}
        sizes = {"train": len(xtr), "test": len(xte)}
    else:
        n_test = max(n_synthetic // 5, 1)
        ds = {"train": synthetic.synthetic_cifar10(n_synthetic, seed), "test": synthetic.synthetic_cifar10(n_test, seed + 1)}
        sizes = {"train": n_synthetic, "test": n_test}
    out = ds if split is None else (ds[split] if isinstance(split, str) else [ds[s] for s in split])
    if with_info:
        info = _ns(name="cifar10", splits={k: _ns(num_examples=v) for k, v in sizes.items()},
                   features=_ns(image_shape=(32, 32, 3), num_classes=10), source="disk" if found else "synthetic")
        return out, info
    return out


def _strategy_ctor(cls):
    """``tf.distribute.MirroredStrategy()`` takes no arguments in the scripts; launched with torchrun (one process per
    GPU, ``WORLD_SIZE`` > 1) the strategy has to join that job instead of standing alone."""
    def ctor(*args, **kw):
        import os
        if "world" not in kw and int(os.environ.get("WORLD_SIZE", "1")) > 1:
            from .parallel.world import World
            kw["world"] = World.from_env(os.environ.get("IDCB_DEVICE"))
        return cls(*args, **kw)
    ctor.__name__ = cls.__name__
    ctor.__doc__ = cls.__doc__
    return ctor


import numpy as _np

tf = _ns(
    float32=_np.float32, double=_np.float64, float64=_np.float64, cast=_cast,
    strings=_ns(split=_strings_split),
    nest=_ns(map_structure=_map_structure),
    compat=_ns(v1=_ns(enable_v2_behavior=lambda: None, py_func=_py_func)),
    keras=_ns(
        Sequential=_apps.Sequential,
        applications=_ns(MobileNetV2=_apps.MobileNetV2, VGG16=_apps.VGG16, DenseNet121=_apps.DenseNet121,
                         DenseNet169=_apps.DenseNet169, DenseNet201=_apps.DenseNet201,
                         vgg16=_ns(VGG16=_apps.VGG16), mobilenet_v2=_ns(MobileNetV2=_apps.MobileNetV2),
                         densenet=_ns(DenseNet121=_apps.DenseNet121, DenseNet169=_apps.DenseNet169,
                                      DenseNet201=_apps.DenseNet201)),
        layers=_ns(Conv2D=_apps.Conv2D, MaxPooling2D=_apps.MaxPooling2D, AveragePooling2D=_apps.AveragePooling2D,
                   Dropout=_apps.Dropout, Flatten=_apps.Flatten, GlobalAveragePooling2D=_apps.GlobalAveragePooling2D,
                   Dense=_apps.Dense),
        models=_ns(Sequential=_apps.Sequential, clone_model=lambda m, **kw: _network.clone_model(
            m.build() if hasattr(m, "build") and not isinstance(m, _network.Network) else m, **kw)),
        optimizers=_ns(RMSprop=_optim.RMSprop, SGD=_optim.SGD, Adam=_optim.Adam),
        losses=_ns(BinaryCrossentropy=_losses.BinaryCrossentropy, CategoricalCrossentropy=_losses.CategoricalCrossentropy),
        metrics=_ns(BinaryAccuracy=_metrics.BinaryAccuracy, CategoricalAccuracy=_metrics.CategoricalAccuracy),
        callbacks=_ns(ModelCheckpoint=_engine.ModelCheckpoint, Callback=_engine.Callback, History=_engine.History,
                      EarlyStopping=_engine.EarlyStopping, LearningRateScheduler=_engine.LearningRateScheduler,
                      ReduceLROnPlateau=_engine.ReduceLROnPlateau, CSVLogger=_engine.CSVLogger),
    ),
    distribute=_ns(MirroredStrategy=_strategy_ctor(_strategy.MirroredStrategy), get_strategy=_strategy.get_strategy,
                   experimental=_ns(CentralStorageStrategy=_strategy_ctor(_strategy.CentralStorageStrategy))),
    data=_ns(Dataset=_dataset.Dataset, experimental=_ns(AUTOTUNE=_dataset.AUTOTUNE)),
    io=_ns(read_file=_read_file),
    image=_ns(decode_png=_decode_png, convert_image_dtype=_convert_image_dtype, resize=_resize),
)

tff = _ns(
    learning=_ns(from_keras_model=_fedavg.from_keras_model,
                 build_federated_averaging_process=_fedavg.build_federated_averaging_process,
                 build_federated_evaluation=_fedavg.build_federated_evaluation,
                 state_with_new_model_weights=_fedavg.state_with_new_model_weights,
                 ModelWeights=_fedavg.ModelWeights),
    simulation=_ns(ClientData=_client_data.ClientData),
)

tfds = _ns(load=_tfds_load)

paillier = _ns(generate_paillier_keypair=_paillier.generate_paillier_keypair, EncryptedNumber=_paillier.EncryptedNumber,
               PaillierPublicKey=_paillier.PaillierPublicKey, PaillierPrivateKey=_paillier.PaillierPrivateKey)

roc_auc_score = _metrics.roc_auc_score
