"""A small host-side dataset pipeline with the ``tf.data`` operations the reference scripts use:
``list_files -> map(process_path) -> take/skip/shard -> cache -> shuffle(1000) -> batch -> prefetch``
(/root/reference/dist_model_tf_mobile.py:34-65,105-113; fed_model.py:65-96,157-165,178-180;
secure_fed_model.py:173-210).

Elements are ``(image float32 [H,W,3] in [0,1], label)`` pairs; batches are
``(torch.float32 [B,H,W,3], torch tensor [B])``. Everything is lazy and re-iterable: each ``iter()``
replays the pipeline (``shuffle`` reshuffles per iteration, like ``reshuffle_each_iteration=True``).
"""
from __future__ import annotations

import glob as _glob
import os
import threading
import queue
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Iterable, Iterator

import numpy as np
import torch

AUTOTUNE = -1


class Dataset:
    def __init__(self, gen_fn: Callable[[], Iterator], length: int | None = None):
        self._gen = gen_fn
        self._len = length

    # ---- sources -------------------------------------------------------------------------------
    @staticmethod
    def list_files(pattern: str, shuffle: bool = True, seed: int | None = None) -> "Dataset":
        files = sorted(_glob.glob(pattern))
        epoch = [0]

        def gen():
            order = list(files)
            if shuffle:
                rng = np.random.default_rng(None if seed is None else seed + epoch[0])
                epoch[0] += 1
                rng.shuffle(order)
            return iter(order)
        return Dataset(gen, len(files))

    @staticmethod
    def from_tensor_slices(tensors) -> "Dataset":
        if isinstance(tensors, tuple):
            n = len(tensors[0])
            return Dataset(lambda: (tuple(t[i] for t in tensors) for i in range(n)), n)
        n = len(tensors)
        return Dataset(lambda: (tensors[i] for i in range(n)), n)

    @staticmethod
    def from_generator(fn: Callable[[], Iterable], length: int | None = None) -> "Dataset":
        return Dataset(lambda: iter(fn()), length)

    # ---- transformations -----------------------------------------------------------------------
    def map(self, fn: Callable, num_parallel_calls: int | None = None) -> "Dataset":
        src = self
        workers = 0 if not num_parallel_calls else (min(8, os.cpu_count() or 1) if num_parallel_calls < 0
                                                    else num_parallel_calls)

        def apply(e):
            return fn(*e) if isinstance(e, tuple) else fn(e)

        batched = getattr(fn, "batched", None)

        def gen():
            if batched is not None and workers > 1:
                # the function brings its own parallel implementation for a list of elements (process_path: one
                # native call decodes a chunk of files on `workers` threads, no Python per file)
                def chunks():
                    chunk = []
                    for e in src:
                        chunk.append(e)
                        if len(chunk) >= 64 * workers:
                            yield from batched(chunk, workers)
                            chunk = []
                    if chunk:
                        yield from batched(chunk, workers)
                return chunks()
            if workers <= 1:
                return (apply(e) for e in src)
            def it():
                with ThreadPoolExecutor(workers) as ex:
                    window = []
                    for e in src:
                        window.append(ex.submit(apply, e))
                        if len(window) >= 4 * workers:
                            yield window.pop(0).result()
                    for f in window:
                        yield f.result()
            return it()
        return Dataset(gen, self._len)

    def take(self, n: int) -> "Dataset":
        src = self

        def gen():
            def it():
                if n <= 0:
                    return
                for i, e in enumerate(src):
                    yield e
                    if i + 1 >= n:
                        break
            return it()
        return Dataset(gen, None if self._len is None else min(n, self._len))

    def skip(self, n: int) -> "Dataset":
        src = self

        def gen():
            def it():
                for i, e in enumerate(src):
                    if i >= n:
                        yield e
            return it()
        return Dataset(gen, None if self._len is None else max(self._len - n, 0))

    def shard(self, num_shards: int, index: int) -> "Dataset":
        src = self

        def gen():
            return (e for i, e in enumerate(src) if i % num_shards == index)
        ln = None if self._len is None else (self._len - index + num_shards - 1) // num_shards
        return Dataset(gen, ln)

    def concatenate(self, other: "Dataset") -> "Dataset":
        src = self

        def gen():
            def it():
                yield from src
                yield from other
            return it()
        ln = None if (self._len is None or other._len is None) else self._len + other._len
        return Dataset(gen, ln)

    def repeat(self, count: int | None = None) -> "Dataset":
        src = self

        def gen():
            def it():
                i = 0
                while count is None or i < count:
                    yield from src
                    i += 1
            return it()
        return Dataset(gen, None if (count is None or self._len is None) else self._len * count)

    def cache(self, filename: str | None = None) -> "Dataset":
        """First full pass materialises the elements in memory (or in ``filename`` via torch.save);
        later passes replay them (dist_model_tf_mobile.py:51-55)."""
        src = self
        store: dict = {"data": None}

        def gen():
            if store["data"] is None and filename and os.path.exists(filename):
                store["data"] = torch.load(filename, weights_only=False)
            if store["data"] is not None:
                return iter(store["data"])

            def it():
                buf = []
                for e in src:
                    buf.append(e)
                    yield e
                store["data"] = buf
                if filename:
                    torch.save(buf, filename)
            return it()
        return Dataset(gen, self._len)

    def shuffle(self, buffer_size: int, seed: int | None = None) -> "Dataset":
        """Streaming buffer shuffle with tf.data's algorithm: keep ``buffer_size`` elements, emit a
        uniformly random one and replace it with the next input."""
        src = self
        epoch = [0]

        def gen():
            rng = np.random.default_rng(None if seed is None else seed + epoch[0])
            epoch[0] += 1

            def it():
                buf = []
                for e in src:
                    if len(buf) < buffer_size:
                        buf.append(e)
                        continue
                    j = int(rng.integers(len(buf)))
                    out, buf[j] = buf[j], e
                    yield out
                while buf:
                    j = int(rng.integers(len(buf)))
                    buf[j], buf[-1] = buf[-1], buf[j]
                    yield buf.pop()
            return it()
        return Dataset(gen, self._len)

    def batch(self, batch_size: int, drop_remainder: bool = False) -> "Dataset":
        src = self

        def collate(items):
            if isinstance(items[0], tuple):
                return tuple(_stack([it[k] for it in items]) for k in range(len(items[0])))
            return _stack(items)

        def gen():
            def it():
                cur = []
                for e in src:
                    cur.append(e)
                    if len(cur) == batch_size:
                        yield collate(cur)
                        cur = []
                if cur and not drop_remainder:
                    yield collate(cur)
            return it()
        ln = None
        if self._len is not None:
            ln = self._len // batch_size if drop_remainder else -(-self._len // batch_size)
        return Dataset(gen, ln)

    def prefetch(self, buffer_size: int = AUTOTUNE) -> "Dataset":
        """Produce elements on a background thread (dist_model_tf_mobile.py:61-63)."""
        src = self
        depth = 2 if buffer_size is None or buffer_size < 0 else max(1, buffer_size)

        def gen():
            q: queue.Queue = queue.Queue(maxsize=depth)
            stop = object()
            err: list = []
            cancelled = threading.Event()

            def put(e) -> bool:
                # a consumer that stopped early (evaluate(steps=20), next(iter(ds))) must not leave this
                # thread blocked on a full queue for ever, with the upstream generators suspended
                while not cancelled.is_set():
                    try:
                        q.put(e, timeout=0.05)
                        return True
                    except queue.Full:
                        continue
                return False

            def work():
                it_src = iter(src)
                try:
                    for e in it_src:
                        if not put(e):
                            break
                except BaseException as ex:  # noqa: BLE001
                    err.append(ex)
                finally:
                    close = getattr(it_src, "close", None)
                    if close is not None:
                        try:
                            close()         # unwinds upstream generators (map's thread pool shuts down)
                        except Exception:   # noqa: BLE001
                            pass
                    put(stop)
            th = threading.Thread(target=work, daemon=True)
            th.start()

            def it():
                try:
                    while True:
                        e = q.get()
                        if e is stop:
                            if err:
                                raise err[0]
                            return
                        yield e
                finally:
                    cancelled.set()         # GeneratorExit / exhaustion / error: release the producer
            return it()
        return Dataset(gen, self._len)

    # ---- consumption ---------------------------------------------------------------------------
    def __iter__(self):
        return self._gen()

    def __len__(self):
        if self._len is None:
            raise TypeError("dataset length is unknown")
        return self._len

    def cardinality(self):
        return self._len


def _stack(items):
    a = items[0]
    if isinstance(a, torch.Tensor):
        return torch.stack(items)
    if isinstance(a, np.ndarray):
        return torch.from_numpy(np.stack(items))
    if isinstance(a, (bool, np.bool_)):
        return torch.tensor(np.asarray(items, dtype=np.float32))
    if isinstance(a, (int, np.integer)):
        return torch.tensor(np.asarray(items, dtype=np.int64))
    if isinstance(a, (float, np.floating)):
        return torch.tensor(np.asarray(items, dtype=np.float32))
    return list(items)


# ---------------------------------------------------------------------------------------------
# File decoding (A2, A3): label = parent dir name == '1'; PNG -> float32 [0,1] -> bilinear resize
# ---------------------------------------------------------------------------------------------
def get_label(file_path: str, as_int: bool = False):
    """dist_model_tf_mobile.py:34-36 (bool) / fed_model.py:65-67 (int)."""
    lab = os.path.basename(os.path.dirname(file_path)) == "1"
    return int(lab) if as_int else lab


_png_lib = None


def _native_png():
    """The host library's PNG stage (csrc/host/png.cpp), or False when the library cannot be loaded."""
    global _png_lib
    if _png_lib is None:
        if os.environ.get("IDCB_NATIVE_PNG", "1") == "0":
            _png_lib = False
            return _png_lib
        try:
            import ctypes as C
            from ..ops import _lib
            lib = _lib.host_lib()
            lib.idcbh_png_decode_f32.argtypes = [C.c_char_p, C.c_longlong, C.c_void_p, C.c_int, C.c_int,
                                                 C.POINTER(C.c_int), C.POINTER(C.c_int)]
            lib.idcbh_png_decode_files.argtypes = [C.POINTER(C.c_char_p), C.c_longlong, C.c_void_p, C.c_int, C.c_int,
                                                   C.c_void_p, C.c_int]
            _png_lib = lib
        except Exception:  # noqa: BLE001 - no compiler / no library: Pillow does the work
            _png_lib = False
    return _png_lib


def _decode_img_pillow(data_or_path, size) -> np.ndarray:
    from PIL import Image
    import io
    img = Image.open(io.BytesIO(data_or_path) if isinstance(data_or_path, (bytes, bytearray)) else data_or_path)
    arr = np.asarray(img.convert("RGB"), dtype=np.float32) / 255.0
    if arr.shape[0] != size[0] or arr.shape[1] != size[1]:
        t = torch.from_numpy(arr).permute(2, 0, 1).unsqueeze(0)
        t = torch.nn.functional.interpolate(t, size=tuple(size), mode="bilinear", align_corners=False)
        arr = t.squeeze(0).permute(1, 2, 0).contiguous().numpy()
    return arr


def decode_img(data_or_path, size=(50, 50)) -> np.ndarray:
    """decode_png(channels=3) -> convert_image_dtype(float32) -> resize bilinear
    (dist_model_tf_mobile.py:37-40). The native decoder (csrc/host/png.cpp: inflate, row filters, colour types,
    /255, bilinear resize; the call releases the GIL, so ``map(num_parallel_calls=...)`` decodes in parallel) handles
    every non-interlaced 8-bit-or-less PNG; anything else, and any file it rejects, goes through Pillow, whose error
    is the one the caller sees for a corrupt file."""
    lib = _native_png()
    if lib:
        try:
            if isinstance(data_or_path, (bytes, bytearray)):
                data = bytes(data_or_path)
            else:
                with open(data_or_path, "rb") as f:
                    data = f.read()
            out = np.empty((int(size[0]), int(size[1]), 3), dtype=np.float32)
            if lib.idcbh_png_decode_f32(data, len(data), out.ctypes.data, int(size[0]), int(size[1]), None, None) == 0:
                return out
        except OSError:
            pass
    return _decode_img_pillow(data_or_path, size)


def decode_files(paths, size=(50, 50), threads: int | None = None, out: np.ndarray | None = None) -> np.ndarray:
    """Decode a list of PNG files into one float32 array [n, H, W, 3] with a pool of native threads (the whole
    ``map(process_path)`` stage of the reference pipeline in one call; files the native decoder leaves alone are
    decoded by Pillow afterwards). ``out`` may be a preallocated (e.g. pinned) array of that shape."""
    paths = [os.fspath(p) for p in paths]
    n, (H, W) = len(paths), (int(size[0]), int(size[1]))
    if out is None:
        out = np.empty((n, H, W, 3), dtype=np.float32)
    if out.shape != (n, H, W, 3) or out.dtype != np.float32 or not out.flags.c_contiguous:
        raise ValueError("decode_files: out must be a C-contiguous float32 array of shape [n, H, W, 3]")
    lib = _native_png()
    todo = range(n)
    if lib and n:
        import ctypes as C
        arr = (C.c_char_p * n)(*[p.encode() for p in paths])
        status = np.zeros(n, dtype=np.uint8)
        threads = threads or min(32, os.cpu_count() or 1)
        lib.idcbh_png_decode_files(arr, n, out.ctypes.data, H, W, status.ctypes.data, int(threads))
        todo = np.nonzero(status)[0].tolist()
    for i in todo:
        out[i] = _decode_img_pillow(paths[i], (H, W))
    return out


def make_process_path(size=(50, 50), label_as_int: bool = False):
    def process_path(file_path: str):
        return decode_img(file_path, size), get_label(file_path, label_as_int)

    def batched(file_paths: list, workers: int):
        """the same mapping for a chunk of paths: one threaded native decode (Dataset.map's fast path)"""
        imgs = decode_files(file_paths, size, threads=workers)
        return [(imgs[i], get_label(p, label_as_int)) for i, p in enumerate(file_paths)]

    if _native_png():
        process_path.batched = batched
    return process_path


def prepare_for_training(ds: Dataset, batch_size: int, cache=True, shuffle_buffer_size: int = 1000,
                         seed: int | None = None) -> Dataset:
    """cache -> shuffle(1000) -> batch -> prefetch (dist_model_tf_mobile.py:47-65)."""
    if cache:
        ds = ds.cache(cache) if isinstance(cache, str) else ds.cache()
    ds = ds.shuffle(buffer_size=shuffle_buffer_size, seed=seed)
    ds = ds.batch(batch_size)
    ds = ds.prefetch(buffer_size=AUTOTUNE)
    return ds
