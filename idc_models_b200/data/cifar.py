"""On-disk CIFAR-10 reader: the data source of the dense script, which the reference gets from
``tfds.load(name='cifar10', as_supervised=True)`` (/root/reference/dist_model_tf_dense.py:120-121).

There is no network here, so nothing is downloaded; a copy the user has placed under ``<path>/data`` is
read in either of the two layouts the dataset is distributed in:

* ``cifar-10-batches-bin``: ``data_batch_{1..5}.bin`` + ``test_batch.bin``, records of 1 label byte followed by
  3072 pixel bytes stored plane by plane (R, G, B; each 32 rows of 32);
* ``cifar-10-batches-py``: ``data_batch_{1..5}`` + ``test_batch``, pickled dicts with ``data`` (N x 3072 uint8,
  same plane order) and ``labels``.

* a ``tensorflow_datasets`` cache, which is what running the reference once leaves behind:
  ``<data_dir>/cifar10/<version>/cifar10-{train,test}.tfrecord-*`` — TFRecord files of ``tf.train.Example`` protos
  with a PNG-encoded ``image`` and an int64 ``label`` (data/tfrecord.py reads the framing and the protos, the native
  PNG decoder the images). Looked for under ``<path>/data``, ``<path>``, ``<path>/tensorflow_datasets``,
  ``$TFDS_DATA_DIR`` and ``~/tensorflow_datasets``.

Images come back as uint8 NHWC arrays (what tfds yields), labels as int64.
"""
from __future__ import annotations

import os
import pickle

import numpy as np

_BIN, _PY = "cifar-10-batches-bin", "cifar-10-batches-py"
_RECORD = 1 + 3 * 32 * 32


def _tfds_dir(root: str) -> str | None:
    """``<root>/cifar10/<version>`` holding train and test TFRecord shards (the newest version if several)"""
    import glob
    base = os.path.join(root, "cifar10")
    if not os.path.isdir(base):
        return None
    for d in sorted((d for d in glob.glob(os.path.join(base, "*")) if os.path.isdir(d)), reverse=True):
        if glob.glob(os.path.join(d, "cifar10-train.tfrecord*")) and glob.glob(os.path.join(d, "cifar10-test.tfrecord*")):
            return d
    return None


def find_cifar10(path: str) -> str | None:
    """Directory of an on-disk CIFAR-10 copy under ``path`` (or ``path/data``), or None: the binary / python batch
    directories first, then a tensorflow_datasets cache (also under $TFDS_DATA_DIR and ~/tensorflow_datasets)."""
    for root in (os.path.join(path, "data"), path):
        for name in (_BIN, _PY):
            for d in (os.path.join(root, name), os.path.join(root, "cifar10", name)):
                if os.path.isdir(d):
                    return d
    roots = [os.path.join(path, "data"), path, os.path.join(path, "tensorflow_datasets")]
    if os.environ.get("TFDS_DATA_DIR"):
        roots.append(os.environ["TFDS_DATA_DIR"])
    roots.append(os.path.join(os.path.expanduser("~"), "tensorflow_datasets"))
    for root in roots:
        d = _tfds_dir(root)
        if d is not None:
            return d
    return None


def _read_tfds(files):
    """``cifar10-<split>.tfrecord-*`` shards -> (uint8 [N, 32, 32, 3], int64 [N])"""
    from . import tfrecord
    from .dataset import decode_img
    xs, ys = [], []
    for f in sorted(files):
        for rec in tfrecord.read_records(f):
            ex = tfrecord.parse_example(rec)
            if not ex.get("image") or not ex.get("label"):
                raise ValueError(f"{f}: a record without image / label features (found {sorted(ex)})")
            img = decode_img(ex["image"][0], (32, 32))                 # float32 in [0, 1], exact multiples of 1/255
            xs.append(np.rint(img * 255.0).astype(np.uint8))
            ys.append(int(ex["label"][0]))
    if not xs:
        raise ValueError("no records in " + ", ".join(files))
    return np.stack(xs), np.asarray(ys, dtype=np.int64)


def _planes_to_nhwc(flat: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(flat.reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1))


def _read_bin(files):
    xs, ys = [], []
    for f in files:
        raw = np.fromfile(f, dtype=np.uint8)
        if raw.size == 0 or raw.size % _RECORD:
            raise ValueError(f"{f}: size {raw.size} is not a whole number of {_RECORD}-byte CIFAR-10 records")
        rec = raw.reshape(-1, _RECORD)
        ys.append(rec[:, 0].astype(np.int64))
        xs.append(_planes_to_nhwc(rec[:, 1:]))
    return np.concatenate(xs), np.concatenate(ys)


class _BatchUnpickler(pickle.Unpickler):
    """The python-version batches are pickles of a dict of bytes / lists / one numpy array. A pickle can name any
    callable, so only the handful of globals such a file needs are resolvable; anything else is refused."""
    _ALLOWED = {("numpy.core.multiarray", "_reconstruct"), ("numpy._core.multiarray", "_reconstruct"),
                ("numpy", "ndarray"), ("numpy", "dtype"), ("copy_reg", "_reconstructor"), ("copyreg", "_reconstructor"),
                ("__builtin__", "object"), ("builtins", "object")}

    def find_class(self, module, name):
        if (module, name) not in self._ALLOWED:
            raise pickle.UnpicklingError(f"CIFAR-10 batch refers to {module}.{name}: not something a dataset file needs")
        return super().find_class(module, name)


def _read_py(files):
    xs, ys = [], []
    for f in files:
        with open(f, "rb") as fh:
            d = _BatchUnpickler(fh, encoding="bytes").load()
        data = d[b"data"] if b"data" in d else d["data"]
        labels = d.get(b"labels", d.get("labels"))
        xs.append(_planes_to_nhwc(np.asarray(data, dtype=np.uint8)))
        ys.append(np.asarray(labels, dtype=np.int64))
    return np.concatenate(xs), np.concatenate(ys)


def load_cifar10(directory: str):
    """((x_train, y_train), (x_test, y_test)) from a ``cifar-10-batches-{bin,py}`` directory or a tfds cache directory."""
    import glob
    shards = glob.glob(os.path.join(directory, "cifar10-train.tfrecord*"))
    if shards:
        return _read_tfds(shards), _read_tfds(glob.glob(os.path.join(directory, "cifar10-test.tfrecord*")))
    is_bin = os.path.basename(os.path.normpath(directory)) == _BIN or os.path.exists(os.path.join(directory, "test_batch.bin"))
    ext = ".bin" if is_bin else ""
    train = [os.path.join(directory, f"data_batch_{i}{ext}") for i in range(1, 6)]
    train = [f for f in train if os.path.exists(f)]
    test = os.path.join(directory, f"test_batch{ext}")
    if not train or not os.path.exists(test):
        raise FileNotFoundError(f"{directory}: expected data_batch_1..5{ext} and test_batch{ext}")
    read = _read_bin if is_bin else _read_py
    return read(train), read([test])
