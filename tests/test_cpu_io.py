"""CPU tests: flat checkpoints and the minimal HDF5 subset behind Keras .h5 import/export."""
import struct

import numpy as np
import pytest

from idc_models_b200.io import checkpoint, keras_h5
from idc_models_b200.models.applications import build_transfer_model


def test_checkpoint_roundtrip_and_resume_rule(tmp_path):
    m = build_transfer_model("mobilenet_v2", device="cpu", seed=1)
    p = tmp_path / "pretrained" / "cp.ckpt"
    assert not checkpoint.exists(str(p))
    m.save_weights(str(p))
    assert checkpoint.exists(str(p)) and p.read_bytes()[:8] == b"IDCBCKP1"
    m2 = build_transfer_model("mobilenet_v2", device="cpu", seed=2)
    m2.load_weights(str(p))
    assert all(np.array_equal(a, b) for a, b in zip(m.get_weights(), m2.get_weights()))
    assert not (tmp_path / "pretrained" / "cp.ckpt.tmp").exists()     # atomic rename


@pytest.mark.parametrize("arch", ["mobilenet_v2", "vgg16", "densenet121"])
def test_keras_h5_roundtrip(arch, tmp_path):
    m = build_transfer_model(arch, device="cpu", seed=3)
    p = str(tmp_path / "w.h5")
    m.save_weights(p)
    data, attrs = keras_h5.read_h5(p)
    base_name = m.base.name
    assert attrs[("", "layer_names")] == [base_name, "global_average_pooling2d", "dense"]
    assert attrs[("", "backend")] == ["tensorflow"]
    wn = attrs[(base_name, "weight_names")]
    assert all(n.endswith(":0") for n in wn) and len(wn) == len(m.weight_names()) - 2
    assert attrs[("global_average_pooling2d", "weight_names")] == []
    # Keras layouts on disk
    if arch == "vgg16":
        assert data["vgg16/block1_conv1/kernel:0"].shape == (3, 3, 3, 64)
    if arch == "mobilenet_v2":
        assert data[base_name + "/expanded_conv_depthwise/depthwise_kernel:0"].shape == (3, 3, 32, 1)
    assert data["dense/dense/kernel:0"].shape[1] == 1
    m2 = build_transfer_model(arch, device="cpu", seed=4)
    m2.load_weights(p)
    assert all(np.array_equal(a, b) for a, b in zip(m.get_weights(), m2.get_weights()))


def test_h5_superblock_and_object_layout_by_hand(tmp_path):
    """byte-level checks of the structures a foreign HDF5 reader depends on"""
    p = str(tmp_path / "t.h5")
    keras_h5.write_h5(p, {"g/sub/x:0": np.arange(6, dtype=np.float32).reshape(2, 3)},
                      [("", "backend", ["tensorflow"], True), ("g", "weight_names", ["sub/x:0"], False)])
    b = open(p, "rb").read()
    assert b[:8] == b"\x89HDF\r\n\x1a\n"
    assert b[8] == 0 and b[13] == 8 and b[14] == 8                       # superblock v0, 8-byte offsets/lengths
    eof, = struct.unpack_from("<Q", b, 40)
    assert eof == len(b)
    root_hdr, cache_type = struct.unpack_from("<Q", b, 64)[0], struct.unpack_from("<I", b, 72)[0]
    assert cache_type == 1 and b[root_hdr] == 1                           # cached symbol-table entry, v1 header
    btree, heap = struct.unpack_from("<QQ", b, 80)
    assert b[btree:btree + 4] == b"TREE" and b[heap:heap + 4] == b"HEAP"
    snod, = struct.unpack_from("<Q", b, btree + 32)
    assert b[snod:snod + 4] == b"SNOD" and struct.unpack_from("<H", b, snod + 6)[0] == 1
    # the payload is stored contiguously, little-endian float32
    raw = np.arange(6, dtype="<f4").tobytes()
    assert raw in b
    data, attrs = keras_h5.read_h5(p)
    assert np.array_equal(data["g/sub/x:0"], np.arange(6, dtype=np.float32).reshape(2, 3))
    assert attrs[("g", "weight_names")] == ["sub/x:0"]


def test_flat_application_style_file_is_accepted(tmp_path):
    """the published *_notop.h5 files have one group per LAYER (no nested model group)"""
    m = build_transfer_model("vgg16", device="cpu", seed=5)
    named = m.named_weights()
    ds, attrs, layer_names = {}, [], []
    for l in m.base.layers:
        if not l.spec.params:
            continue
        layer_names.append(l.name)
        attrs.append((l.name, "weight_names", [p.name + ":0" for p in l.spec.params], False))
        for p_ in l.spec.params:
            ds[f"{l.name}/{p_.name}:0"] = named[p_.name]
    attrs.append(("", "layer_names", layer_names, False))
    p = str(tmp_path / "vgg16_notop.h5")
    keras_h5.write_h5(p, ds, attrs)
    m2 = build_transfer_model("vgg16", device="cpu", seed=6)
    missing = keras_h5.load_into(m2, p, strict=False)
    assert sorted(missing) == ["dense/bias", "dense/kernel"]
    assert np.array_equal(m2.get_weights()[0], m.get_weights()[0])
    with pytest.raises(KeyError):
        keras_h5.load_into(m2, p, strict=True)


def test_not_an_hdf5_file_is_a_clear_error(tmp_path):
    p = tmp_path / "x.h5"
    p.write_bytes(b"hello world" * 100)
    with pytest.raises(IOError, match="not an HDF5 file"):
        keras_h5.read_h5(str(p))


def test_full_model_h5_with_optimizer_state_roundtrip(tmp_path):
    """``model.save('m.h5')``: weights under /model_weights, JSON configs, optimizer slots under /optimizer_weights;
    a fresh model loads the weights from it and its optimizer resumes from the saved state."""
    import json
    import torch
    from idc_models_b200.io import keras_h5
    from idc_models_b200.models.applications import build_transfer_model
    from idc_models_b200.optim.optimizers import RMSprop

    def make(seed):
        m = build_transfer_model("mobilenet_v2_tiny", (50, 50, 3), 1, device="cpu", seed=seed)
        m.base.trainable = True
        m.compile(optimizer=RMSprop(lr=1e-3), loss="bce", metrics=["accuracy"])
        return m
    g = torch.Generator().manual_seed(0)
    x, y = torch.rand(8, 50, 50, 3, generator=g), (torch.rand(8, generator=g) > 0.5).float()
    a = make(0)
    for _ in range(2):
        a.train_on_batch(x, y)
    path = str(tmp_path / "model.h5")
    a.save(path)
    data, attrs = keras_h5.read_h5(path)
    assert any(k.startswith("model_weights/") and k.endswith("/kernel:0") for k in data)
    assert "optimizer_weights/RMSprop/iter:0" in data and float(data["optimizer_weights/RMSprop/iter:0"][0]) == 2.0
    assert any(k.startswith("optimizer_weights/RMSprop/") and k.endswith("/rms:0") for k in data)
    cfg = [v for (grp, name), v in attrs.items() if name == "model_config"] if isinstance(attrs, dict) else None
    if cfg:
        assert json.loads(cfg[0][0] if isinstance(cfg[0], (list, tuple)) else cfg[0])["class_name"] == "Sequential"
    b = make(1)
    b.load_weights(path)
    assert keras_h5.load_optimizer_state(b, path) > 0
    assert torch.equal(a.store.master, b.store.master) and torch.equal(a.store.stats, b.store.stats)
    assert b.optimizer.iterations == 2 and torch.allclose(a.optimizer.s1, b.optimizer.s1)
    # both continue identically
    a.train_on_batch(x, y)
    b.train_on_batch(x, y)
    assert torch.allclose(a.store.master, b.store.master, atol=1e-7)


# ---------------------------------------------------------------------------------------------------------
# A file laid out the way the HDF5 library writes "new style" files (libver='latest'): version-2 superblock,
# version-2 object headers ("OHDR", with a continuation chunk "OCHK"), compact groups made of link messages,
# version-3 attributes with Keras's chunked name lists (weight_names0 / weight_names1). Built byte by byte here —
# nothing of it comes from this repository's writer, which only emits version-0/1 structures.
# ---------------------------------------------------------------------------------------------------------
import struct

UNDEF = 0xFFFFFFFFFFFFFFFF


def _msg(t, body, flags=0):
    return struct.pack("<BHB", t, len(body), flags) + body


def _ohdr(messages, chunk_pad=0):
    """version-2 object header, no times, 2-byte chunk-0 size field"""
    body = b"".join(messages) + b"\0" * chunk_pad
    return b"OHDR" + bytes([2, 0x01]) + struct.pack("<H", len(body)) + body + b"\0\0\0\0"   # (checksum not verified)


def _f32_type():
    return bytes([0x11, 0x20, 0x1F, 0x00]) + struct.pack("<I", 4) + struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)


def _space(*dims):
    return bytes([2, len(dims), 0, 1]) + b"".join(struct.pack("<Q", d) for d in dims)


def _str_attr(name, values):
    width = max(len(v) for v in values) + 1
    dt = bytes([0x13, 0x00, 0x00, 0x00]) + struct.pack("<I", width)
    ds = _space(len(values))
    nm = name.encode() + b"\0"
    data = b"".join(v.encode().ljust(width, b"\0") for v in values)
    return _msg(0x0C, bytes([3, 0]) + struct.pack("<HHH", len(nm), len(dt), len(ds)) + b"\0" + nm + dt + ds + data)


def _link(name, addr):
    return _msg(0x06, bytes([1, 0x00, len(name)]) + name.encode() + struct.pack("<Q", addr))


def _link_info(heap=UNDEF):
    return _msg(0x02, bytes([0, 0]) + struct.pack("<QQ", heap, UNDEF))


def _build_v2_file(dense_root=False):
    kernel = np.arange(12, dtype=np.float32).reshape(3, 4) / 7
    bias = np.array([0.5, -1.25, 2.0, 3.5], dtype=np.float32)
    blob = bytearray(b"\0" * 48)                                     # superblock goes here

    def put(b):
        while len(blob) % 8:
            blob.append(0)
        at = len(blob)
        blob.extend(b)
        return at

    def dataset(arr):
        data_at = put(arr.tobytes())
        return put(_ohdr([_msg(0x01, _space(*arr.shape)), _msg(0x03, _f32_type()),
                          _msg(0x08, bytes([3, 1]) + struct.pack("<QQ", data_at, arr.nbytes))]))

    k_at, b_at = dataset(kernel), dataset(bias)
    # the "dense" group: its link messages live in a continuation chunk
    cont_msgs = _link("kernel:0", k_at) + _link("bias:0", b_at)
    cont_at = put(b"OCHK" + cont_msgs + b"\0\0\0\0")
    inner = put(_ohdr([_link_info(), _msg(0x10, struct.pack("<QQ", cont_at, len(cont_msgs) + 8))]))
    # Keras nests: /dense (top-level layer group) / dense (name scope) / kernel:0
    top = put(_ohdr([_link_info(), _link("dense", inner),
                     _str_attr("weight_names0", ["dense/kernel:0"]), _str_attr("weight_names1", ["dense/bias:0"])]))
    root = put(_ohdr([_link_info(0x1234 if dense_root else UNDEF), _link("dense", top),
                      _str_attr("layer_names", ["dense"]), _str_attr("backend", ["tensorflow"])]))
    sb = bytes([0x89]) + b"HDF\r\n\x1a\n" + bytes([2, 8, 8, 0]) + struct.pack("<QQQQ", 0, UNDEF, len(blob), root) + b"\0\0\0\0"
    blob[:len(sb)] = sb
    return bytes(blob), kernel, bias


def test_reader_handles_version2_headers_link_messages_and_chunked_name_attributes(tmp_path):
    from idc_models_b200.io import keras_h5
    raw, kernel, bias = _build_v2_file()
    path = tmp_path / "new_style.h5"
    path.write_bytes(raw)
    data, attrs = keras_h5.read_h5(str(path))
    assert set(data) == {"dense/dense/kernel:0", "dense/dense/bias:0"}
    assert np.array_equal(data["dense/dense/kernel:0"], kernel) and np.array_equal(data["dense/dense/bias:0"], bias)
    assert attrs[("", "layer_names")] == ["dense"]
    assert attrs[("dense", "weight_names0")] == ["dense/kernel:0"] and attrs[("dense", "weight_names1")] == ["dense/bias:0"]

    # and the name resolution used by load_weights finds both tensors in it
    class Net:
        def weight_names(self):
            return ["dense/kernel", "dense/bias"]

        def set_named_weights(self, found, strict=False):
            self.found = found
            return [n for n in self.weight_names() if n not in found]

    net = Net()
    assert keras_h5.load_into(net, str(path)) == []
    assert np.array_equal(net.found["dense/kernel"], kernel)


def test_reader_names_the_unsupported_feature_of_a_dense_storage_group(tmp_path):
    """groups with many members are stored in a fractal heap + v2 B-tree by new-style files: not in the supported
    subset, and the error has to say so instead of returning an empty file"""
    from idc_models_b200.io import keras_h5
    raw, _, _ = _build_v2_file(dense_root=True)
    path = tmp_path / "dense_links.h5"
    path.write_bytes(raw)
    with pytest.raises(IOError, match="dense link storage"):
        keras_h5.read_h5(str(path))


def test_h5_and_checkpoint_readers_under_sanitizers(tmp_path):
    """tests/native/fuzz_h5.cpp: truncations, bit flips, saturated bytes and random damage of a file from our writer
    and of the hand-built version-2 file; every mutation is opened, listed and every listed dataset read, with the
    readers compiled under -fsanitize=address,undefined. A length or offset taken from the file and
    trusted shows up as a sanitizer abort. (This harness found the unchecked header / name / attribute reads that the
    checked accessors of csrc/host/hdf5.cpp now guard.)"""
    import shutil
    import subprocess
    from pathlib import Path
    from idc_models_b200.io import keras_h5
    root = Path(__file__).resolve().parent.parent
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    exe = tmp_path / "fuzz_h5"
    host = root / "csrc" / "host"
    cmd = [cxx, "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-pthread",
           "-I", str(root / "include"), "-I", str(host), str(host / "hdf5.cpp"), str(host / "ckpt.cpp"),
           str(host / "host_runtime.cpp"), str(root / "tests" / "native" / "fuzz_h5.cpp"), "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and any(t in r.stderr.lower() for t in ("sanitize", "asan", "ubsan", "tsan")):
        pytest.skip("this g++ has no sanitizer runtime")
    assert r.returncode == 0, r.stderr[-3000:]
    rng = np.random.default_rng(0)
    ours = tmp_path / "w.h5"
    keras_h5.write_h5(str(ours), {"conv1/conv1/kernel:0": rng.normal(size=(3, 3, 3, 8)).astype(np.float32),
                                  "conv1/conv1/bias:0": np.zeros(8, np.float32), "bn/bn/gamma:0": np.ones(8, np.float32)},
                      [("/", "layer_names", ["conv1", "bn"], False), ("/", "backend", ["tensorflow"], True),
                       ("conv1", "weight_names", ["conv1/kernel:0", "conv1/bias:0"], False),
                       ("bn", "weight_names", ["bn/gamma:0"], False)])
    blob = _build_v2_file()
    v2 = tmp_path / "v2.h5"
    v2.write_bytes(blob[0] if isinstance(blob, tuple) else blob)
    # a checkpoint container too (the harness pushes every mutated file through both readers)
    import ctypes as C
    from idc_models_b200.ops import _lib
    hl = _lib.host_lib()
    hl.idcbh_ckpt_begin.restype = C.c_void_p
    ck = tmp_path / "w.ckpt"
    h = C.c_void_p(hl.idcbh_ckpt_begin(str(ck).encode()))
    w = rng.normal(size=(4, 6)).astype(np.float32)
    assert hl.idcbh_ckpt_add_f32(h, b"dense/kernel", w.ctypes.data_as(C.c_void_p), (C.c_longlong * 2)(4, 6), 2) == 0
    assert hl.idcbh_ckpt_end(h) == 0
    r = subprocess.run([str(exe), "3", str(tmp_path / "mut.h5"), str(ours), str(v2), str(ck)], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-300:], r.stderr[-3000:])
    files, opened, reads = [int(t) for t in r.stdout.split() if t.isdigit()]
    assert files > 3000 and 0 < opened < files and reads > opened


def test_checkpoint_with_a_lying_header_is_refused(tmp_path):
    """a header length larger than the file, and a tensor whose declared byte count disagrees with its shape, are
    errors (they used to size a read)"""
    import ctypes as C
    import struct
    from idc_models_b200.io import checkpoint
    from idc_models_b200.ops import _lib
    hl = _lib.host_lib()
    hl.idcbh_ckpt_begin.restype = C.c_void_p
    ck = tmp_path / "w.ckpt"
    h = C.c_void_p(hl.idcbh_ckpt_begin(str(ck).encode()))
    w = np.arange(24, dtype=np.float32).reshape(4, 6)
    assert hl.idcbh_ckpt_add_f32(h, b"dense/kernel", w.ctypes.data_as(C.c_void_p), (C.c_longlong * 2)(4, 6), 2) == 0
    assert hl.idcbh_ckpt_end(h) == 0
    good = ck.read_bytes()
    assert np.array_equal(checkpoint.read_checkpoint(str(ck))["dense/kernel"], w)
    for lie in (0xFFFFFFFFFFFFFFFF, 0x7FFFFFFFFFFFFFFF, len(good)):
        bad = tmp_path / "bad.ckpt"
        bad.write_bytes(good[:8] + struct.pack("<Q", lie) + good[16:])
        with pytest.raises(IOError, match="exceeds the file"):
            checkpoint.read_checkpoint(str(bad))
    (hdr_len,) = struct.unpack("<Q", good[8:16])
    header = good[16:16 + hdr_len].decode()
    assert '"nbytes"' in header
    lying = header.replace('"nbytes": 96', '"nbytes": 960').replace('"nbytes":96', '"nbytes":960')
    assert lying != header
    pad = lying.encode()
    bad = tmp_path / "bad2.ckpt"
    bad.write_bytes(good[:8] + struct.pack("<Q", len(pad)) + pad + good[16 + hdr_len:])
    with pytest.raises(IOError, match="declares"):
        checkpoint.read_checkpoint(str(bad))
