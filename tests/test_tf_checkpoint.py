"""The TensorBundle (TensorFlow V2 checkpoint) reader / writer of obst/run/tf_checkpoint.py, without TensorFlow:
published known answers for the pieces (CRC-32C check values, crc masking, the table magic, protobuf wire bytes
assembled by hand from tensor_bundle.proto) and the round trip.  CPU only (the checksum is host code of the library)."""
import os
import struct
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from obst import inputs, native  # noqa: E402
from obst.run import tf_checkpoint as tfc  # noqa: E402


def test_native_crc32c_against_known_answers_and_the_bytewise_statement():
    assert native.crc32c(b"123456789") == 0xE3069283                    # the CRC-32C check value
    assert native.crc32c(bytes(32)) == 0x8A9136AA                       # RFC 3720 B.4
    assert native.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43
    assert native.crc32c(bytes(range(32))) == 0x46DD794E
    assert native.crc32c(b"") == 0
    rng = np.random.default_rng(0)
    lib = native.load()
    for n in (1, 7, 8, 9, 4095, 4096, 12287, 12288, 12289, 40000):
        a = rng.integers(0, 256, n, dtype=np.uint8)
        want = inputs.crc32c_py(a.tobytes())
        assert native.crc32c(a) == want == lib.obst_crc32c_portable(a.ctypes.data, n, 0), n
        k = n // 3
        assert native.crc32c(a[k:].copy(), native.crc32c(a[:k].copy())) == want   # continuation
    t = torch.from_numpy(rng.standard_normal(1000).astype(np.float32))
    assert native.crc32c(t) == inputs.crc32c_py(t.numpy().tobytes())


def test_crc_mask_round_trips_and_matches_the_tfrecord_mask():
    for c in (0, 1, 0xE3069283, 0xFFFFFFFF, 0x12345678):
        assert tfc.unmask_crc(tfc.mask_crc(c)) == c
    assert tfc.mask_crc(0) == 0xA282EAD8
    assert tfc.mask_crc(native.crc32c(b"abc")) == inputs.masked_crc32c(b"abc")


def test_protobuf_wire_bytes():
    # BundleHeaderProto{num_shards: 1, version{producer: 1}}: 08 01 | 1a 02 (08 01)
    assert tfc.encode_header(1) == bytes.fromhex("0801" "1a02" "0801")
    assert tfc.decode_header(tfc.encode_header(3)) == {"num_shards": 3, "endianness": 0, "producer": 1, "min_consumer": 0}
    # BundleEntryProto{dtype: DT_FLOAT, shape{dim{size: 2} dim{size: 3}}, size: 24, crc32c: 0x12345678}
    want = bytes.fromhex("0801" "1208" "12020802" "12020803" "2818" "35" "78563412")
    assert tfc.encode_entry(tfc.DT_FLOAT, [2, 3], 0, 0, 24, 0x12345678) == want
    e = tfc.decode_entry(want)
    assert (e["dtype"], e["shape"], e["shard_id"], e["offset"], e["size"], e["crc32c"]) == (1, [2, 3], 0, 0, 24, 0x12345678)
    # a scalar: the (empty) shape message is still written; large offsets are multi-byte varints
    scalar = tfc.encode_entry(tfc.DT_INT64, [], 0, 300, 8, 1)
    assert scalar == bytes.fromhex("0809" "1200" "20ac02" "2808" "35" "01000000")
    assert tfc.decode_entry(scalar)["shape"] == [] and tfc.decode_entry(scalar)["offset"] == 300
    # a zero-sized dim is an empty Dim message
    assert tfc.decode_entry(tfc.encode_entry(tfc.DT_FLOAT, [0, 4], 0, 0, 0, 0))["shape"] == [0, 4]
    assert tfc.put_varint(-1) == bytes([0xFF] * 9 + [0x01])
    assert tfc.get_varint(tfc.put_varint(2 ** 40 + 5), 0) == (2 ** 40 + 5, 6)


def test_table_layout_and_round_trip():
    items = [(b"", b"header")] + [(f"var/{i:05d}/weights".encode(), os.urandom(20 + i % 7)) for i in range(1000)]
    items.sort()
    data = tfc.build_table(items, block_size=4096)     # many blocks, each with several restart points
    assert data[-8:] == bytes.fromhex("57fb808b247547db")          # kTableMagicNumber, little-endian
    assert tfc.read_table(data) == items
    one = tfc.build_table(items)                                    # default block size: one data block
    assert tfc.read_table(one) == items and len(one) < len(data)
    # prefix compression inside a block: second key shares "abc" with the first
    blk = tfc._BlockBuilder(16)
    blk.add(b"abc1", b"x")
    blk.add(b"abc2", b"yz")
    raw = blk.finish()
    assert raw == b"\x00\x04\x01abc1x" + b"\x03\x01\x022yz" + struct.pack("<II", 0, 1)
    # a restart point every 16 entries
    blk = tfc._BlockBuilder(16)
    for i in range(40):
        blk.add(f"k{i:03d}".encode(), b"v")
    assert struct.unpack("<I", blk.finish()[-4:])[0] == 3
    # corruption is detected by the block checksum
    bad = bytearray(data)
    bad[100] ^= 1
    with pytest.raises(ValueError, match="checksum"):
        tfc.read_table(bytes(bad))
    with pytest.raises(ValueError, match="magic"):
        tfc.read_table(data[:-1] + b"\x00")
    with pytest.raises(ValueError, match="increasing"):
        tfc.build_table([(b"b", b""), (b"a", b"")])


def test_bundle_round_trip(tmp_path):
    rng = np.random.default_rng(1)
    tensors = {
        "gpt0v/body0v/0_0v/bottleneck_group_linear_0v/get_variable0": torch.from_numpy(rng.standard_normal((8, 16, 4)).astype(np.float32)),
        "gpt0v/body0v/0_0v/bottleneck_group_linear_0v/get_variable0/sm3_momentum_0.9_1_1/dim0": torch.rand(8),
        "bf16/weights": torch.from_numpy(rng.standard_normal((5, 7)).astype(np.float32)).to(torch.bfloat16),
        "global_step": torch.tensor(1502, dtype=torch.int64),
        "empty": torch.zeros((0, 3)),
        "ints": torch.arange(12, dtype=torch.int32).view(3, 4),
    }
    prefix = str(tmp_path / "run" / "model.ckpt-1502")
    tfc.write_bundle(prefix, tensors)
    assert sorted(os.listdir(tmp_path / "run")) == ["model.ckpt-1502.data-00000-of-00001", "model.ckpt-1502.index"]
    size = sum(t.numel() * t.element_size() for t in tensors.values())
    assert os.path.getsize(prefix + ".data-00000-of-00001") == size          # raw bytes, back to back
    back = tfc.read_bundle(prefix)
    assert back.keys() == tensors.keys()
    for k, t in tensors.items():
        assert back[k].dtype == t.dtype and back[k].shape == t.shape and torch.equal(back[k], t), k
    assert tfc.read_bundle(prefix, names={"global_step"}).keys() == {"global_step"}
    # the data file holds the tensors in key order at the offsets the index states
    with open(prefix + ".index", "rb") as f:
        items = tfc.read_table(f.read())
    assert items[0][0] == b"" and [k for k, _ in items] == sorted(k for k, _ in items)
    e = tfc.decode_entry(dict(items)[b"ints"])
    raw = open(prefix + ".data-00000-of-00001", "rb").read()[e["offset"]:e["offset"] + e["size"]]
    assert np.array_equal(np.frombuffer(raw, dtype="<i4"), np.arange(12))
    assert tfc.unmask_crc(e["crc32c"]) == native.crc32c(raw)
    # a flipped data byte is caught by the tensor checksum
    with open(prefix + ".data-00000-of-00001", "r+b") as f:
        f.seek(e["offset"])
        f.write(b"\xff")
    with pytest.raises(ValueError, match="checksum"):
        tfc.read_bundle(prefix)
    assert "ints" in tfc.read_bundle(prefix, verify=False)


def test_checkpoint_state_file(tmp_path):
    tfc.write_checkpoint_state(str(tmp_path), "model.ckpt-20", ["model.ckpt-10"])
    text = open(tmp_path / "checkpoint").read()
    assert text == ('model_checkpoint_path: "model.ckpt-20"\nall_model_checkpoint_paths: "model.ckpt-10"\n'
                    'all_model_checkpoint_paths: "model.ckpt-20"\n')
    assert tfc.latest_checkpoint(str(tmp_path)) == str(tmp_path / "model.ckpt-20")
    assert tfc.latest_checkpoint(str(tmp_path / "nowhere")) is None
