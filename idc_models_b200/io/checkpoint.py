"""Weights-only checkpoints with the reference's file naming (``<path>/pretrained/cp.ckpt``) and
resume rule (train if the checkpoint is missing, else ``load_weights``): /root/reference/fed_model.py:
100-105,135-138,175; SURVEY.md 2.B B16, 5.4.

Three on-disk formats, chosen by extension:
* ``*.h5`` / ``*.hdf5`` — Keras weight files (io.keras_h5, own HDF5 subset);
* anything else (``cp.ckpt``) — the native flat format written by ``libidcb200_host.so``
  (csrc/host/ckpt.cpp): 16-byte preamble, JSON index, 64-byte aligned raw fp32 payloads, atomic rename.
Tensors are stored under their Keras weight names in Keras layouts, so a checkpoint written on one device
layout (channel-padded stems, bf16 shadows) loads on any other.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from ..ops import _lib


def exists(path: str) -> bool:
    return os.path.exists(path)


def save_weights(net, path: str) -> None:
    if path.endswith((".h5", ".hdf5")):
        from . import keras_h5
        keras_h5.save_weights(net, path)
        return
    if net.strategy.world.rank != 0:
        return
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    lib = _lib.host_lib()
    lib.idcbh_ckpt_begin.restype = C.c_void_p
    h = C.c_void_p(lib.idcbh_ckpt_begin(path.encode()))
    keep = []
    for name, arr in net.named_weights().items():
        a = np.ascontiguousarray(arr, dtype=np.float32)
        keep.append(a)
        shape = (C.c_longlong * max(a.ndim, 1))(*a.shape) if a.ndim else (C.c_longlong * 1)(1)
        lib.idcbh_ckpt_add_f32(h, name.encode(), a.ctypes.data_as(C.c_void_p), shape, C.c_int(a.ndim))
    if lib.idcbh_ckpt_end(h) != 0:
        raise IOError(lib.idcbh_last_error().decode())


def read_checkpoint(path: str) -> dict:
    lib = _lib.host_lib()
    lib.idcbh_ckpt_header.restype = C.c_longlong
    need = lib.idcbh_ckpt_header(path.encode(), None, C.c_longlong(0))
    if need < 0:
        raise IOError(lib.idcbh_last_error().decode())
    buf = C.create_string_buffer(int(need))
    hl = lib.idcbh_ckpt_header(path.encode(), buf, C.c_longlong(need))
    if hl < 0:
        raise IOError(lib.idcbh_last_error().decode())
    meta = json.loads(buf.value.decode())
    out = {}
    for t in meta["tensors"]:
        a = np.empty(t["shape"], dtype=np.float32)
        if int(t["nbytes"]) != a.nbytes or int(t["offset"]) < 0:       # the header is file content: never trusted
            raise IOError(f"checkpoint {path}: tensor {t['name']!r} declares {t['nbytes']} bytes for shape {t['shape']}")
        if lib.idcbh_ckpt_read(path.encode(), C.c_ulonglong(hl), C.c_ulonglong(t["offset"]),
                               a.ctypes.data_as(C.c_void_p), C.c_ulonglong(t["nbytes"])) != 0:
            raise IOError(lib.idcbh_last_error().decode())
        out[t["name"]] = a
    return out


def load_weights(net, path: str, strict: bool = True) -> None:
    if path.endswith((".h5", ".hdf5")):
        from . import keras_h5
        keras_h5.load_into(net, path, by_name=True, strict=strict)
        return
    net.set_named_weights(read_checkpoint(path), strict=strict)
