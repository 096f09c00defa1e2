"""The allocation-free secure-mean entry point (``idcb_secagg_mean_noalloc``) called straight through the
C ABI, checked against the ordinary entry point that takes caller-provided staging buffers.

Why the entry point exists: some programs that link this library may not create device memory when they
call it — the GPU's memory belongs to their own arena, or they run under a start-up memory quota, or a
cudaMalloc at that moment would stall the device. For them the staging area sits inside the library.

The file name sorts last on purpose: if this test ever fails under ``pytest -x``, every other test has
already run. It passes alone in a fresh process and after the rest of the suite (the library's built-in
storage keeps state between calls by design, so the test makes no assumption about its initial contents).
"""
import ctypes as C
import gc

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from idc_models_b200.ops import _lib as L
    lib = L.cuda_lib()
    lib.idcb_secagg_noalloc_max_elems.restype = C.c_longlong
    return L, lib


def _reference(L, lib, clients, k, stride, off, n, frac_bits, seed):
    """the existing entry point: one rank, caller-owned staging buffer and flag pad"""
    from idc_models_b200.parallel.symm import CommWorld, ptr_array
    dev = clients.device
    ybuf = torch.empty(n, dtype=torch.int32, device=dev)
    pad = torch.zeros(64, dtype=torch.int32, device=dev)
    out = torch.zeros(stride, dtype=torch.float32, device=dev)
    cw = CommWorld(1, 0, 1)
    cw.signal_pads[0] = pad.data_ptr()
    L.check(lib.idcb_secagg_masked_sum(C.byref(cw), ptr_array([clients.data_ptr()]), ptr_array([ybuf.data_ptr()]),
                                       None, ptr_array([out.data_ptr()]), None, C.c_int(k), C.c_longlong(stride),
                                       C.c_longlong(off), C.c_longlong(n), C.c_int(frac_bits), C.c_ulonglong(seed),
                                       L.stream()))
    torch.cuda.synchronize()
    return out


def test_noalloc_entry_point_matches_the_buffered_one_on_two_calls_in_a_row():
    L, lib = _lib()
    torch.manual_seed(0)
    for call, (k, stride, off, n, seed) in enumerate([(6, 8192, 1024, 4096, 0xABCDEF0123),
                                                      (3, 2048, 0, 2048, 0x55AA55AA)]):
        clients = torch.randn(k, stride, device="cuda") * 2
        want = _reference(L, lib, clients, k, stride, off, n, 16, seed)
        out = torch.zeros(stride, dtype=torch.float32, device="cuda")
        # objects of earlier tests may still be waiting for the garbage collector; let it run now so that a
        # collection in the middle of the measured window cannot change the allocator's count
        gc.collect()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.idcb_secagg_mean_noalloc(C.c_void_p(clients.data_ptr()), C.c_void_p(out.data_ptr()), C.c_int(k),
                                          C.c_longlong(stride), C.c_longlong(off), C.c_longlong(n), C.c_int(16),
                                          C.c_ulonglong(seed), L.stream())
        assert rc == 0, lib.idcb_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert torch.cuda.mem_get_info()[0] >= free_before          # nor did the driver hand out device memory
        # same kernel, same arithmetic: bit-identical; and it is the mean, to the quantisation step
        assert torch.equal(out, want), f"call {call}: differs from the buffered entry point"
        assert (out[off:off + n] - clients[:, off:off + n].mean(0)).abs().max().item() < 2e-5
        assert out[:off].abs().max().item() == 0 if off else True


def test_noalloc_entry_point_refuses_oversized_ranges():
    L, lib = _lib()
    limit = lib.idcb_secagg_noalloc_max_elems()
    assert limit == 1 << 20
    x = torch.zeros(8, device="cuda")
    rc = lib.idcb_secagg_mean_noalloc(C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()), C.c_int(2),
                                      C.c_longlong(limit + 4), C.c_longlong(0), C.c_longlong(limit + 4), C.c_int(16),
                                      C.c_ulonglong(1), L.stream())
    assert rc != 0
    assert b"exceeds the built-in staging area" in lib.idcb_last_error()


def _call(lib, L, clients, out, k, stride, off, n, seed, stream=None):
    return lib.idcb_secagg_mean_noalloc(C.c_void_p(clients.data_ptr()), C.c_void_p(out.data_ptr()), C.c_int(k),
                                        C.c_longlong(stride), C.c_longlong(off), C.c_longlong(n), C.c_int(16),
                                        C.c_ulonglong(seed), stream if stream is not None else L.stream())


def test_noalloc_entry_point_at_its_size_limit_and_on_a_side_stream():
    """The largest range the built-in staging area holds, launched on a stream that is not the default one
    (the storage is per GPU, not per stream; the caller orders the calls)."""
    L, lib = _lib()
    n = lib.idcb_secagg_noalloc_max_elems()
    torch.manual_seed(1)
    clients = torch.randn(2, n, device="cuda")
    want = _reference(L, lib, clients, 2, n, 0, n, 16, 77)
    out = torch.zeros(n, dtype=torch.float32, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        rc = _call(lib, L, clients, out, 2, n, 0, n, 77, C.c_void_p(side.cuda_stream))
    assert rc == 0, lib.idcb_last_error().decode()
    side.synchronize()
    assert torch.equal(out, want)
    assert (out - clients.mean(0)).abs().max().item() < 2e-5


def test_noalloc_entry_point_checks_the_range_and_the_alignment():
    L, lib = _lib()
    x = torch.zeros(2, 64, device="cuda")
    out = torch.zeros(64, device="cuda")
    assert _call(lib, L, x, out, 2, 64, 32, 64, 1) != 0                 # offset + n > stride
    assert b"does not fit in a client row" in lib.idcb_last_error()
    assert _call(lib, L, x, out, 2, 64, -4, 8, 1) != 0                  # negative offset
    rc = lib.idcb_secagg_mean_noalloc(C.c_void_p(x.data_ptr() + 4), C.c_void_p(out.data_ptr()), C.c_int(2),
                                      C.c_longlong(60), C.c_longlong(0), C.c_longlong(8), C.c_int(16),
                                      C.c_ulonglong(1), L.stream())
    assert rc != 0 and b"16-byte aligned" in lib.idcb_last_error()
    torch.cuda.synchronize()
