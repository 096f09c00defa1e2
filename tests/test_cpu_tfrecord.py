"""TFRecord / tf.train.Example reader (data/tfrecord.py) and the tensorflow_datasets CIFAR-10 cache layout
(data/cifar.py): what `tfds.load('cifar10')` of /root/reference/dist_model_tf_dense.py:120 leaves on a user's disk."""
import io
import os
import struct

import numpy as np
import pytest

from idc_models_b200.data import cifar, tfrecord as T

Image = pytest.importorskip("PIL.Image")


def test_crc32c_and_example_wire_format_against_literals():
    assert T.crc32c(b"123456789") == 0xE3069283           # the CRC-32C check value
    assert T.crc32c(b"") == 0 and T.crc32c(b"\x00" * 32) == 0x8A9136AA
    # tf.train.Example(features=Features(feature={'label': Feature(int64_list=Int64List(value=[7]))})) serialised
    literal = b"\n\x10\n\x0e\n\x05label\x12\x05\x1a\x03\n\x01\x07"
    assert T.build_example({"label": 7}) == literal
    assert T.parse_example(literal) == {"label": [7]}
    # an unpacked int64 list (one varint per value, as older writers emit) and a negative value
    unpacked = b"\n\x16\n\x14\n\x01x\x12\x0f\x1a\x0d\x08\x03\x08" + b"\xff" * 9 + b"\x01"
    assert T.parse_example(unpacked) == {"x": [3, -1]}
    ex = T.parse_example(T.build_example({"image": b"\x89PNG", "label": 3, "id": b"train_1", "w": [0.5, -2.0]}))
    assert ex == {"image": [b"\x89PNG"], "label": [3], "id": [b"train_1"], "w": [0.5, -2.0]}


def test_record_framing_round_trip_and_damage(tmp_path):
    p = tmp_path / "a.tfrecord"
    payloads = [b"", b"x", os.urandom(1000), T.build_example({"label": 1})]
    T.write_records(str(p), payloads)
    assert list(T.read_records(str(p), verify=True)) == payloads
    raw = bytearray(p.read_bytes())
    # first record: length 0 -> 8 zero bytes, then the masked CRC of those 8 bytes
    assert raw[:8] == b"\x00" * 8 and struct.unpack("<I", raw[8:12])[0] == T.masked_crc(b"\x00" * 8)
    bad = bytearray(raw)
    bad[2] ^= 1                                            # a length byte
    (tmp_path / "b.tfrecord").write_bytes(bad)
    with pytest.raises(IOError, match="length checksum"):
        list(T.read_records(str(tmp_path / "b.tfrecord")))
    bad = bytearray(raw)
    bad[-10] ^= 1                                          # inside the last payload
    (tmp_path / "c.tfrecord").write_bytes(bad)
    assert len(list(T.read_records(str(tmp_path / "c.tfrecord")))) == 4         # not verified by default
    with pytest.raises(IOError, match="data checksum"):
        list(T.read_records(str(tmp_path / "c.tfrecord"), verify=True))
    (tmp_path / "d.tfrecord").write_bytes(raw[:-3])
    with pytest.raises(IOError, match="truncated"):
        list(T.read_records(str(tmp_path / "d.tfrecord")))


def _png(a):
    b = io.BytesIO()
    Image.fromarray(a, "RGB").save(b, "PNG")
    return b.getvalue()


def test_tfds_cifar10_cache_is_found_and_read(tmp_path, monkeypatch):
    rng = np.random.default_rng(0)
    d = tmp_path / "data" / "cifar10" / "3.0.2"
    d.mkdir(parents=True)
    xs = {"train": rng.integers(0, 256, (12, 32, 32, 3), dtype=np.uint8), "test": rng.integers(0, 256, (5, 32, 32, 3), dtype=np.uint8)}
    ys = {"train": rng.integers(0, 10, 12), "test": rng.integers(0, 10, 5)}
    for split, shards in (("train", 2), ("test", 1)):
        n = len(xs[split])
        for s in range(shards):
            idx = range(s, n, shards)
            T.write_records(str(d / f"cifar10-{split}.tfrecord-{s:05d}-of-{shards:05d}"),
                            [T.build_example({"image": _png(xs[split][i]), "label": int(ys[split][i]),
                                              "id": f"{split}_{i:05d}".encode()}) for i in idx])
    (d / "dataset_info.json").write_text("{}")
    monkeypatch.setenv("HOME", str(tmp_path / "nohome"))
    found = cifar.find_cifar10(str(tmp_path))
    assert found == str(d)
    (xtr, ytr), (xte, yte) = cifar.load_cifar10(found)
    order = list(range(0, 12, 2)) + list(range(1, 12, 2))                        # shard 0 then shard 1
    assert xtr.dtype == np.uint8 and np.array_equal(xtr, xs["train"][order]) and np.array_equal(ytr, ys["train"][order])
    assert np.array_equal(xte, xs["test"]) and np.array_equal(yte, ys["test"])
    # the same cache through the reference's call, and through $TFDS_DATA_DIR
    from idc_models_b200.compat import tfds
    datasets, info = tfds.load(name="cifar10", with_info=True, as_supervised=True, data_dir=str(tmp_path / "data"))
    assert info.source == "disk" and info.splits["train"].num_examples == 12
    image, label = next(iter(datasets["test"]))
    assert np.array_equal(np.asarray(image), xs["test"][0]) and int(label) == int(ys["test"][0])
    monkeypatch.setenv("TFDS_DATA_DIR", str(tmp_path / "data"))
    assert cifar.find_cifar10(str(tmp_path / "elsewhere")) == str(d)
