"""Native PNG stage of the input pipeline (csrc/host/png.cpp) against Pillow: every colour type / bit depth /
compression level Pillow can write, the row filters real images exercise, packed palettes, the fallback for the kinds
the native decoder leaves alone (16-bit, interlaced), corrupt files, the bilinear resize against the PyTorch
statement, and the threaded file-list decode (/root/reference/dist_model_tf_mobile.py:37-45 is the stage replaced)."""
import ctypes as C
import io

import numpy as np
import pytest

from idc_models_b200.data import dataset as D

Image = pytest.importorskip("PIL.Image")


def _lib():
    lib = D._native_png()
    if not lib:
        pytest.skip("host library not available")
    return lib


def _native(lib, data, h, w):
    out = np.empty((h, w, 3), np.float32)
    sw, sh = C.c_int(), C.c_int()
    rc = lib.idcbh_png_decode_f32(data, len(data), out.ctypes.data, h, w, C.byref(sw), C.byref(sh))
    return rc, out, (sw.value, sh.value)


def _pil(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"), dtype=np.float32) / 255.0


def _smooth(rng, h, w, c):
    y, x = np.mgrid[0:h, 0:w]
    a = np.stack([(np.sin(x / 7 + i) + np.cos(y / 5 - i)) * 60 + 128 + rng.normal(0, 6, (h, w)) for i in range(c)], -1)
    return np.clip(a, 0, 255).astype(np.uint8)


def _png(im, **kw):
    b = io.BytesIO()
    im.save(b, "PNG", **kw)
    return b.getvalue()


@pytest.mark.parametrize("hw", [(50, 50), (1, 1), (3, 7), (64, 33), (120, 201)])
def test_native_png_is_bit_identical_to_pillow(hw):
    lib = _lib()
    rng = np.random.default_rng(hw[0] * 1000 + hw[1])
    h, w = hw
    for mode in ("RGB", "RGBA", "L", "LA", "P", "1"):
        for lvl in (0, 1, 6, 9):          # 0: stored blocks; 1 with noise: fixed / short dynamic codes; 6, 9: dynamic
            if mode in ("RGB", "RGBA", "L", "LA"):
                c = {"RGB": 3, "RGBA": 4, "L": 1, "LA": 2}[mode]
                arr = _smooth(rng, h, w, c) if lvl != 1 else rng.integers(0, 256, (h, w, c), dtype=np.uint8)
                im = Image.fromarray(arr[:, :, 0] if c == 1 else arr, mode)
            elif mode == "P":
                im = Image.fromarray(_smooth(rng, h, w, 3), "RGB").quantize([2, 4, 16, 200][lvl % 4] if lvl else 256)
            else:
                im = Image.fromarray(rng.random((h, w)) > 0.5).convert("1")
            data = _png(im, compress_level=lvl)
            rc, out, size = _native(lib, data, h, w)
            assert rc == 0, (mode, lvl, lib.idcbh_last_error())
            assert size == (w, h)
            assert np.array_equal(out, _pil(data)), (mode, lvl)


def test_packed_palettes_and_fallback_kinds():
    lib = _lib()
    rng = np.random.default_rng(3)
    for bits in (1, 2, 4):
        data = _png(Image.fromarray(_smooth(rng, 37, 53, 3), "RGB").quantize(2 ** bits), bits=bits)
        rc, out, _ = _native(lib, data, 37, 53)
        assert rc == 0 and np.array_equal(out, _pil(data)), bits
    # 16-bit samples: left to the general decoder (status 1), and decode_img then goes through Pillow
    data16 = _png(Image.fromarray(rng.integers(0, 65535, (20, 20)).astype(np.uint16)))
    rc, _, _ = _native(lib, data16, 20, 20)
    assert rc == 1 and b"16-bit" in lib.idcbh_last_error()
    assert D.decode_img(data16, (20, 20)).shape == (20, 20, 3)
    # a damaged stream and a truncated file are refused with a reason, never decoded into garbage
    good = bytearray(_png(Image.fromarray(_smooth(rng, 20, 20, 3), "RGB")))
    bad = bytearray(good)
    bad[60] ^= 0xFF
    assert _native(lib, bytes(bad), 20, 20)[0] == 2 and lib.idcbh_last_error() != b""
    assert _native(lib, bytes(good[:50]), 20, 20)[0] == 2
    assert _native(lib, b"not a png at all, not even close........", 20, 20)[0] == 2
    flipped = bytearray(good)
    flipped[-20] ^= 0x01                 # inside the adler32 / last IDAT bytes
    assert _native(lib, bytes(flipped), 20, 20)[0] == 2
    with pytest.raises(Exception):
        D.decode_img(bytes(good[:50]), (20, 20))          # Pillow's error surfaces for a corrupt file


@pytest.mark.parametrize("size", [(50, 50), (28, 28), (10, 10), (100, 120)])
def test_native_resize_matches_the_pytorch_statement(size):
    _lib()
    rng = np.random.default_rng(5)
    data = _png(Image.fromarray(_smooth(rng, 73, 91, 3), "RGB"))
    got = D.decode_img(data, size)
    want = D._decode_img_pillow(data, size)
    assert got.shape == want.shape == (*size, 3)
    assert np.abs(got - want).max() < 1e-6


def test_decode_files_threaded_with_fallback(tmp_path):
    _lib()
    rng = np.random.default_rng(9)
    paths = []
    for i in range(40):
        p = tmp_path / f"{i:03d}.png"
        if i == 7:
            Image.fromarray(rng.integers(0, 65535, (50, 50)).astype(np.uint16)).save(p)      # native status 1 -> Pillow
        elif i % 3 == 0:
            Image.fromarray(_smooth(rng, 61, 47, 3), "RGB").save(p)                         # resized
        else:
            Image.fromarray(_smooth(rng, 50, 50, 3), "RGB").save(p)
        paths.append(str(p))
    out = D.decode_files(paths, (50, 50), threads=4)
    assert out.shape == (40, 50, 50, 3)
    for i, p in enumerate(paths):
        assert np.abs(out[i] - D._decode_img_pillow(p, (50, 50))).max() < 1e-6, i
    one = D.decode_files(paths[:1], (50, 50), threads=8)
    assert np.array_equal(one[0], out[0])
    assert D.decode_files([], (50, 50)).shape == (0, 50, 50, 3)


def test_map_fast_path_equals_per_file_mapping(tmp_path):
    """Dataset.map(process_path, num_parallel_calls=AUTOTUNE) takes the chunked native path and yields exactly the
    (image, label) pairs, in order, of the per-file mapping"""
    _lib()
    rng = np.random.default_rng(11)
    for cls in ("0", "1"):
        (tmp_path / cls).mkdir()
    for i in range(150):
        Image.fromarray(_smooth(rng, 50, 50, 3), "RGB").save(tmp_path / str(i % 2) / f"{i:03d}.png")
    files = D.Dataset.list_files(str(tmp_path / "*" / "*.png"), shuffle=False)
    pp = D.make_process_path((50, 50), label_as_int=True)
    assert hasattr(pp, "batched")
    fast = list(files.map(pp, num_parallel_calls=D.AUTOTUNE))
    slow = list(files.map(pp))
    assert len(fast) == len(slow) == 150
    for (a, la), (b, lb) in zip(fast, slow):
        assert la == lb and np.array_equal(a, b)
    assert sorted({l for _, l in fast}) == [0, 1]
