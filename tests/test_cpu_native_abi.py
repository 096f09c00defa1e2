"""The CUDA library on a box without a GPU: it loads (static cudart, no device needed), what it exports is exactly
what include/idcb200.h declares (and likewise for the host library and include/idcb200_host.h), both headers are
valid C99 and C++17, and the entry points check their arguments BEFORE they touch the device — a bad call is a
non-zero return with a message, never a crash; a good call on this box fails cleanly with the CUDA error."""
import ctypes as C
import re
import shutil
import subprocess
from pathlib import Path

import pytest

from idc_models_b200.ops import _lib
from idc_models_b200.parallel.symm import CommWorld

ROOT = Path(__file__).resolve().parent.parent
vp = C.c_void_p


def _exports(lib: Path, prefix: str) -> set:
    nm = shutil.which("nm")
    if nm is None:
        pytest.skip("no nm")
    out = subprocess.run([nm, "-D", "--defined-only", str(lib)], capture_output=True, text=True, check=True).stdout
    return {l.split()[-1] for l in out.splitlines() if l.split() and l.split()[-1].startswith(prefix)}


def _declared(header: Path, prefix: str) -> set:
    return set(re.findall(r"\b(" + prefix + r"[a-z0-9_]*)\s*\(", header.read_text()))


def test_exports_match_the_headers():
    _lib.cuda_lib(); _lib.host_lib()
    native = ROOT / "idc_models_b200" / "_native"
    assert _exports(native / "libidcb200.so", "idcb_") == _declared(ROOT / "include" / "idcb200.h", "idcb_")
    assert _exports(native / "libidcb200_host.so", "idcbh_") == _declared(ROOT / "include" / "idcb200_host.h", "idcbh_")


@pytest.mark.parametrize("header", ["idcb200.h", "idcb200_host.h"])
def test_headers_are_valid_c_and_cpp(header, tmp_path):
    cc, cxx = shutil.which("gcc"), shutil.which("g++")
    if cc is None or cxx is None:
        pytest.skip("no compiler")
    src = tmp_path / "t.c"
    src.write_text(f'#include "{header}"\nint main(void) {{ return 0; }}\n')
    for cmd in ([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror"], [cxx, "-x", "c++", "-std=c++17", "-Wall", "-Werror"]):
        r = subprocess.run([*cmd, "-fsyntax-only", "-I", str(ROOT / "include"), str(src)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr


def test_entry_points_validate_before_touching_the_device():
    lib = _lib.cuda_lib()
    err = lambda: lib.idcb_last_error().decode()
    cw = CommWorld(1, 0, 1)
    cw.signal_pads[0] = 0x1000
    P = (vp * 8)()
    P[0] = 0x2000
    offs, lens = (C.c_longlong * 3)(128, 64, 8000), (C.c_longlong * 3)(1920, 2048, 192)
    args = lambda n_ranges, base, out: (C.byref(cw), P, P, P, P, None, None, C.c_int(4), C.c_longlong(8192), C.c_int(n_ranges),
                                       offs, lens, C.c_float(1.0), vp(base), vp(out), C.c_longlong(8192), None)
    assert lib.idcb_fedavg_round(*args(3, 0x3000, 0x4000)) != 0 and "ascending" in err()
    assert lib.idcb_fedavg_round(*args(17, 0x3000, 0x4000)) != 0 and "at most 16 ranges" in err()
    assert lib.idcb_fedavg_round(*args(1, 0x3000, 0x3000)) != 0 and "distinct" in err()
    assert lib.idcb_fedavg_round(*args(1, 0x3004, 0x4000)) != 0 and "aligned" in err()
    # small CNN: the indexed form needs an index list at least as wide as the batch
    o6 = (C.c_int * 6)(0, 864, 896, 4992, 5000, 5008)
    cnn = lambda idx, stride: lib.idcb_small_cnn_step_indexed(
        vp(0x1000), vp(0x1000), vp(idx), C.c_longlong(stride), C.c_int(37), vp(0x1000), vp(0x1000), vp(0x1000), vp(0x1000),
        C.c_int(4), C.c_int(16), C.c_int(6000), o6, C.c_int(3), C.c_int(1), C.c_float(0.1), C.c_float(0.25), C.c_float(0.5),
        C.c_ulonglong(1), C.c_uint(1), C.c_float(0.0), C.c_int(0), None)
    assert cnn(0, 37) != 0 and "index" in err()
    assert cnn(0x1000, 8) != 0 and "index" in err()
    # secure sum: the range has to fit in a client row, pointers have to be 16-byte aligned
    sec = lambda off, n, ptr: lib.idcb_secagg_masked_sum(C.byref(cw), (vp * 8)(ptr), (vp * 8)(0x2000), None, (vp * 8)(0x3000), None,
                                                         C.c_int(2), C.c_longlong(1024), C.c_longlong(off), C.c_longlong(n),
                                                         C.c_int(16), C.c_ulonglong(1), None)
    assert sec(512, 1024, 0x1000) != 0 and "does not fit" in err()
    assert sec(0, 1024, 0x1004) != 0 and "aligned" in err()
    assert sec(0, 1022, 0x1000) != 0 and "multiples of 4" in err()
    # the allocation-free entry point states its limit
    assert lib.idcb_secagg_noalloc_max_elems() == 1 << 20
    # a well-formed call on a box without a GPU: a CUDA error with a message, not a crash
    # (only where there is no device: with one, the made-up pointers would reach a kernel)
    import torch
    if not torch.cuda.is_available():
        rc = lib.idcb_optim_step(C.c_int(1), vp(0x1000), vp(0x2000), vp(0x3000), vp(0), vp(0), C.c_longlong(1024),
                                 C.c_float(1e-3), C.c_float(0.9), C.c_float(0.0), C.c_float(1e-7), C.c_float(1.0), C.c_int(1), None)
        assert rc != 0 and "CUDA error" in err()
