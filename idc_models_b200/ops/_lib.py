"""ctypes binding to the in-tree native libraries (``_native/libidcb200.so`` and
``_native/libidcb200_host.so``); declarations mirror ``include/idcb200.h``.

The CUDA library is mandatory on a GPU box: if a CUDA device is visible and the library cannot be
loaded, :func:`cuda_lib` raises instead of letting callers fall back to eager PyTorch — a silent
fallback would make every GPU test pass without running a single one of our kernels.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_NATIVE = Path(__file__).resolve().parent.parent / "_native"
_cuda = None
_host = None

c_float_p = C.POINTER(C.c_float)
vp = C.c_void_p


class ConvDesc(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "batch", "h", "w", "cin", "ho", "wo", "cout", "kh", "kw", "stride", "pad_t", "pad_l",
        "x_pitch", "x_coff", "y_pitch", "y_coff", "w_pitch", "w_off")]


class ConvFused(C.Structure):
    _fields_ = [
        ("pro_scale", vp), ("pro_shift", vp), ("pro_act", C.c_int),
        ("epi_scale", vp), ("epi_shift", vp), ("epi_act", C.c_int),
        ("aux", vp), ("aux_pitch", C.c_int), ("aux_coff", C.c_int),
        ("stat_a", vp), ("stat_b", vp), ("bn", vp),
    ]


class BnInline(C.Structure):
    _fields_ = [
        ("sum", vp), ("sq", vp), ("gamma", vp), ("beta", vp),
        ("count", C.c_float), ("eps", C.c_float), ("momentum", C.c_float),
        ("scale_out", vp), ("shift_out", vp), ("mean_out", vp), ("rstd_out", vp),
        ("moving_mean", vp), ("moving_var", vp),
    ]


class BnTail(C.Structure):
    _fields_ = [
        ("out", vp), ("out_pitch", C.c_int), ("out_coff", C.c_int), ("mean", vp), ("rstd", vp),
        ("count", C.c_float), ("dgamma", vp), ("dbeta", vp), ("counter", vp), ("frozen", C.c_int),
    ]


class SliceTail(C.Structure):
    _fields_ = [
        ("out", vp), ("out_pitch", C.c_int), ("out_coff", C.c_int), ("acc_s1", vp), ("acc_s2", vp),
        ("mean", vp), ("rstd", vp), ("bn_mean", vp), ("bn_rstd", vp), ("dgamma", vp), ("dbeta", vp),
        ("count", C.c_float), ("counter", vp), ("c0", C.c_int),
    ]


class ConvBwdFused(C.Structure):
    _fields_ = [
        ("mode", C.c_int), ("aux", vp), ("aux_pitch", C.c_int), ("aux_coff", C.c_int),
        ("bwd_scale", vp), ("bwd_shift", vp), ("bwd_act", C.c_int),
        ("gacc", vp), ("g_pitch", C.c_int), ("g_coff", C.c_int),
        ("stat_a", vp), ("stat_b", vp), ("tail", vp), ("slice", vp),
    ]


class NativeError(RuntimeError):
    pass


def _build_if_missing(path: Path, which: str) -> None:
    if path.exists() or os.environ.get("IDCB_NO_AUTOBUILD"):
        return
    from .. import build as _b
    (_b.build_cuda if which == "cuda" else _b.build_host)()


def cuda_lib() -> C.CDLL:
    """The CUDA kernel library. Raises NativeError if it cannot be loaded."""
    global _cuda
    if _cuda is None:
        path = _NATIVE / "libidcb200.so"
        try:
            _build_if_missing(path, "cuda")
            lib = C.CDLL(str(path))
        except Exception as e:  # noqa: BLE001
            raise NativeError(
                f"idc_models_b200: cannot load {path} ({e}). Run `python -m idc_models_b200.build`; "
                "the GPU path never falls back to eager PyTorch.") from e
        lib.idcb_last_error.restype = C.c_char_p
        _cuda = lib
    return _cuda


def host_lib() -> C.CDLL:
    """The host-only C++ library (HDF5 subset, checkpoint writer, prefetch ring)."""
    global _host
    if _host is None:
        path = _NATIVE / "libidcb200_host.so"
        try:
            _build_if_missing(path, "host")
            lib = C.CDLL(str(path))
        except Exception as e:  # noqa: BLE001
            raise NativeError(f"idc_models_b200: cannot load {path} ({e})") from e
        lib.idcbh_last_error.restype = C.c_char_p
        _host = lib
    return _host


def check(rc: int) -> None:
    if rc != 0:
        msg = cuda_lib().idcb_last_error().decode(errors="replace")
        raise NativeError(f"idcb200 native call failed ({rc}): {msg}")


def ptr(t: torch.Tensor | None) -> vp:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return vp(0)
    return vp(t.data_ptr())


def stream() -> vp:
    return vp(torch.cuda.current_stream().cuda_stream)


_launches = 0


def count_launch(n: int = 1) -> None:
    """Book-keeping for bench.py's ``gpu_launches`` claim: every Python wrapper that launches one of
    our kernels bumps this counter (CUDA-graph replays add the captured count, see train.graph)."""
    global _launches
    _launches += n


def launches() -> int:
    return _launches


def reset_launches() -> None:
    global _launches
    _launches = 0
