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


def _chunk(tag: bytes, body: bytes) -> bytes:
    import struct
    import zlib
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def _filter_rows(img: np.ndarray, filters) -> bytes:
    """PNG-filter an [h, w, c] uint8 image with the given per-row filter types (the encoder side of the five filters,
    written from the specification)"""
    h, w, c = img.shape
    rows = img.reshape(h, w * c).astype(np.int32)
    out = bytearray()
    prev = np.zeros(w * c, np.int32)
    for y in range(h):
        cur = rows[y]
        left = np.concatenate([np.zeros(c, np.int32), cur[:-c]])
        upleft = np.concatenate([np.zeros(c, np.int32), prev[:-c]])
        ft = filters[y % len(filters)]
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - left
        elif ft == 2:
            f = cur - prev
        elif ft == 3:
            f = cur - ((left + prev) >> 1)
        else:
            p = left + prev - upleft
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
            f = cur - pred
        out.append(ft)
        out += (f & 0xFF).astype(np.uint8).tobytes()
        prev = cur
    return bytes(out)


@pytest.mark.parametrize("strategy", ["default", "fixed", "rle", "huffman_only", "stored"])
def test_hand_built_pngs_every_filter_every_deflate_block_type(strategy):
    """PNG files assembled by hand: each of the five row filters forced, zlib told to emit fixed-Huffman, RLE,
    Huffman-only, stored and ordinary dynamic blocks, the compressed stream cut into many IDAT chunks, long
    back-references (image larger than the 32 KiB window)"""
    import struct
    import zlib
    lib = _lib()
    rng = np.random.default_rng(17)
    strat = {"default": zlib.Z_DEFAULT_STRATEGY, "fixed": zlib.Z_FIXED, "rle": zlib.Z_RLE,
             "huffman_only": zlib.Z_HUFFMAN_ONLY, "stored": zlib.Z_DEFAULT_STRATEGY}[strategy]
    level = 0 if strategy == "stored" else 6
    for (h, w, c), filters in [((40, 31, 3), [0]), ((40, 31, 3), [1]), ((40, 31, 3), [2]), ((40, 31, 3), [3]),
                               ((40, 31, 3), [4]), ((130, 150, 3), [4, 3, 2, 1, 0]), ((64, 64, 4), [1, 4]),
                               ((33, 17, 1), [3, 4, 0]), ((200, 200, 3), [0, 4])]:
        img = _smooth(rng, h, w, c)
        if (h, w) == (200, 200):
            img[100:] = img[:100]                     # matches at distances up to the full window
        co = zlib.compressobj(level, zlib.DEFLATED, 15, 9, strat)
        z = co.compress(_filter_rows(img, filters)) + co.flush()
        ctype = {1: 0, 3: 2, 4: 6}[c]
        png = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0))
        png += _chunk(b"tEXt", b"Comment\0hand built")
        step = max(1, len(z) // 7)
        for i in range(0, len(z), step):                   # several IDAT chunks, boundaries anywhere in the stream
            png += _chunk(b"IDAT", z[i:i + step])
        png += _chunk(b"IEND", b"")
        rc, out, size = _native(lib, png, h, w)
        assert rc == 0, (strategy, (h, w, c), filters, lib.idcbh_last_error())
        want = img[:, :, :3] if c >= 3 else np.repeat(img, 3, axis=2)
        assert size == (w, h) and np.array_equal(out, want.astype(np.float32) / 255.0), (strategy, (h, w, c), filters)
        assert np.array_equal(out, _pil(png))              # and Pillow reads the hand-built file the same way


def test_truncations_and_bit_flips_never_crash_or_decode_silently():
    """every prefix of a valid file and a sweep of single-bit flips inside the compressed stream: the decoder either
    rejects the file (status 2) or — when the flip hits a byte that does not matter (padding after the final block) or
    still yields a stream whose adler32 checks — returns exactly what Pillow returns; it never reads out of bounds"""
    lib = _lib()
    rng = np.random.default_rng(23)
    data = _png(Image.fromarray(_smooth(rng, 24, 24, 3), "RGB"))
    full = _pil(data)
    n_rejected = 0
    for cut in range(0, len(data) - 1, 3):
        rc, out, _ = _native(lib, data[:cut] if cut else b"\0", 24, 24)
        assert rc in (0, 2), cut
        if rc == 0:                        # only the trailer (IEND, or the last chunk's CRC) is missing
            assert cut > data.index(b"IEND") - 8 and np.array_equal(out, full), cut
        n_rejected += rc == 2
    assert n_rejected > len(data) // 3 - 10
    idat = data.index(b"IDAT")
    for pos in range(idat + 4, len(data) - 16):
        for bit in (0, 3, 7):
            bad = bytearray(data)
            bad[pos] ^= 1 << bit
            rc, out, _ = _native(lib, bytes(bad), 24, 24)
            assert rc in (0, 2)
            if rc == 0:
                assert np.array_equal(out, _pil(data)), (pos, bit)


def test_decoder_under_address_and_undefined_behaviour_sanitizers(tmp_path):
    """tests/native/fuzz_png.cpp: prefixes, bit flips and random damage of four files, from exact-size heap copies,
    with the decoder compiled under -fsanitize=address,undefined (what a sanitizer can check on the
    host side of this code base). Any out-of-bounds access or undefined shift aborts the harness."""
    import shutil
    import subprocess
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    exe = tmp_path / "fuzz_png"
    cmd = [cxx, "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-pthread",
           "-I", str(root / "include"), "-I", str(root / "csrc" / "host"), str(root / "csrc" / "host" / "png.cpp"),
           str(root / "csrc" / "host" / "host_runtime.cpp"), str(root / "tests" / "native" / "fuzz_png.cpp"), "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and any(t in r.stderr.lower() for t in ("sanitize", "asan", "ubsan", "tsan")):
        pytest.skip("this g++ has no sanitizer runtime")
    assert r.returncode == 0, r.stderr[-3000:]
    rng = np.random.default_rng(1)
    files = []
    for i, mode in enumerate(("RGB", "RGBA", "L", "P")):
        a = _smooth(rng, 40, 37, 4)
        im = {"RGB": lambda: Image.fromarray(a[:, :, :3], "RGB"), "RGBA": lambda: Image.fromarray(a, "RGBA"),
              "L": lambda: Image.fromarray(a[:, :, 0], "L"),
              "P": lambda: Image.fromarray(a[:, :, :3], "RGB").quantize(16)}[mode]()
        p = tmp_path / f"f{i}.png"
        im.save(p)
        files.append(str(p))
    r = subprocess.run([str(exe), "3", *files], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-3000:])
    calls, decoded = [int(t) for t in r.stdout.split() if t.isdigit()]
    assert calls > 5000 and 0 < decoded < calls // 10
