"""TFRecord files of ``tf.train.Example`` protos — the on-disk form of a ``tensorflow_datasets`` cache, which is what a
user of the reference has after ``tfds.load('cifar10')`` has run once (/root/reference/dist_model_tf_dense.py:120:
``~/tensorflow_datasets/cifar10/<version>/cifar10-{train,test}.tfrecord-0000x-of-0000y``).

Two small formats, both written out here from their specifications (no TensorFlow, no protobuf package):

* TFRecord framing: ``uint64 length | uint32 masked_crc32c(length) | data | uint32 masked_crc32c(data)``, little
  endian; the CRC is CRC-32C (Castagnoli), masked as ``((crc >> 15 | crc << 17) + 0xa282ead8) mod 2^32``;
* ``tf.train.Example``: ``Example{features=1: Features{feature=1: map<string, Feature>}}``,
  ``Feature{bytes_list=1 | float_list=2 | int64_list=3}``, each list a ``repeated value = 1`` (packed or not).

A writer is included for the tests (and for exporting a dataset in a form TensorFlow tools read).
"""
from __future__ import annotations

import struct
from typing import Iterator

import numpy as np

# ---- CRC-32C -----------------------------------------------------------------------------------------------
_POLY = 0x82F63B78
_TABLE = []
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ (_POLY if _c & 1 else 0)
    _TABLE.append(_c)
_TABLE_NP = np.array(_TABLE, dtype=np.uint32)


def crc32c(data: bytes) -> int:
    c = 0xFFFFFFFF
    t = _TABLE
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ---- TFRecord framing ----------------------------------------------------------------------------------------
def read_records(path: str, verify: bool = False) -> Iterator[bytes]:
    """Yield the payload of every record. ``verify`` checks both CRCs (pure Python: slow on large files; the length
    CRC is always checked, which is what catches a file that is not a TFRecord or is cut short)."""
    with open(path, "rb") as f:
        while True:
            head = f.read(12)
            if not head:
                return
            if len(head) < 12:
                raise IOError(f"{path}: truncated record header")
            (n,), (crc,) = struct.unpack("<Q", head[:8]), struct.unpack("<I", head[8:])
            if masked_crc(head[:8]) != crc:
                raise IOError(f"{path}: length checksum mismatch (not a TFRecord file, or damaged)")
            data = f.read(n)
            tail = f.read(4)
            if len(data) < n or len(tail) < 4:
                raise IOError(f"{path}: truncated record")
            if verify and masked_crc(data) != struct.unpack("<I", tail)[0]:
                raise IOError(f"{path}: data checksum mismatch")
            yield data


def write_records(path: str, payloads) -> None:
    with open(path, "wb") as f:
        for d in payloads:
            head = struct.pack("<Q", len(d))
            f.write(head + struct.pack("<I", masked_crc(head)) + d + struct.pack("<I", masked_crc(d)))


# ---- protobuf wire format (the three wire types Example uses) ---------------------------------------------------
def _varint(buf: bytes, p: int) -> tuple[int, int]:
    v = s = 0
    while True:
        if p >= len(buf):
            raise ValueError("truncated varint")
        b = buf[p]
        p += 1
        v |= (b & 0x7F) << s
        if b < 0x80:
            return v, p
        s += 7
        if s > 70:
            raise ValueError("varint too long")


def _fields(buf: bytes) -> Iterator[tuple[int, int, object]]:
    """(field number, wire type, value) — value is an int (varint, fixed32/64 raw) or a bytes slice (length-delimited)"""
    p, n = 0, len(buf)
    while p < n:
        key, p = _varint(buf, p)
        num, wt = key >> 3, key & 7
        if wt == 0:
            v, p = _varint(buf, p)
        elif wt == 1:
            v, p = buf[p:p + 8], p + 8
        elif wt == 2:
            ln, p = _varint(buf, p)
            if p + ln > n:
                raise ValueError("truncated length-delimited field")
            v, p = buf[p:p + ln], p + ln
        elif wt == 5:
            v, p = buf[p:p + 4], p + 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield num, wt, v


def _signed64(v: int) -> int:
    return v - (1 << 64) if v >= 1 << 63 else v


def parse_example(buf: bytes) -> dict:
    """{feature name: list of bytes | list of float | list of int}"""
    out: dict = {}
    for num, wt, features in _fields(buf):
        if num != 1 or wt != 2:
            continue
        for fnum, fwt, entry in _fields(features):            # map<string, Feature> entries
            if fnum != 1 or fwt != 2:
                continue
            name, feat = None, b""
            for enum_, ewt, v in _fields(entry):
                if enum_ == 1 and ewt == 2:
                    name = bytes(v).decode("utf-8", "replace")
                elif enum_ == 2 and ewt == 2:
                    feat = v
            if name is None:
                continue
            values: list = []
            for knum, kwt, lst in _fields(feat):
                if kwt != 2:
                    continue
                if knum == 1:                                      # BytesList
                    values = [bytes(v) for n_, w_, v in _fields(lst) if n_ == 1 and w_ == 2]
                elif knum == 2:                                    # FloatList (packed or one fixed32 per value)
                    for n_, w_, v in _fields(lst):
                        if n_ == 1 and w_ == 2:
                            values.extend(struct.unpack(f"<{len(v) // 4}f", v))
                        elif n_ == 1 and w_ == 5:
                            values.append(struct.unpack("<f", v)[0])
                elif knum == 3:                                    # Int64List (packed or one varint per value)
                    for n_, w_, v in _fields(lst):
                        if n_ == 1 and w_ == 2:
                            q = 0
                            while q < len(v):
                                x, q = _varint(v, q)
                                values.append(_signed64(x))
                        elif n_ == 1 and w_ == 0:
                            values.append(_signed64(v))
            out[name] = values
    return out


def _enc_varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _ld(num: int, payload: bytes) -> bytes:
    return _enc_varint(num << 3 | 2) + _enc_varint(len(payload)) + payload


def build_example(features: dict) -> bytes:
    """The inverse of :func:`parse_example` (values: bytes / float / int, or lists of them)."""
    entries = b""
    for name in sorted(features):
        vals = features[name]
        vals = vals if isinstance(vals, (list, tuple)) else [vals]
        if all(isinstance(v, (bytes, bytearray)) for v in vals):
            feat = _ld(1, b"".join(_ld(1, bytes(v)) for v in vals))
        elif all(isinstance(v, (int, np.integer)) for v in vals):
            feat = _ld(3, _ld(1, b"".join(_enc_varint(int(v)) for v in vals)))
        else:
            feat = _ld(2, _ld(1, struct.pack(f"<{len(vals)}f", *[float(v) for v in vals])))
        entries += _ld(1, _ld(1, name.encode()) + _ld(2, feat))
    return _ld(1, entries)
