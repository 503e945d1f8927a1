"""obst.inputs: the text branch of src/inputs.py without TensorFlow -- TFRecord framing, Example wire format, the two
decoders, per-record windowing, file sharding, interleave, batching and the x/y split (src/inputs.py:15-30,
231-268, 528-568)."""
import os
import struct

import numpy as np
import pytest

from tests.helpers import orc  # noqa: F401  (puts the repo on sys.path the way the other tests do)

from obst import inputs
from obst.configs import mixer_config
from obst.dataclass import ModelParameter


def test_crc32c_known_answers():
    assert inputs.crc32c(b"123456789") == 0xE3069283            # the CRC-32C check value
    assert inputs.crc32c(b"") == 0 and inputs.crc32c(bytes(32)) == 0x8A9136AA   # RFC 3720 B.4: 32 zero bytes
    c = inputs.crc32c(b"abc")
    assert inputs.masked_crc32c(b"abc") == ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def test_tfrecord_round_trip_and_corruption(tmp_path):
    path = str(tmp_path / "a.tfrecord")
    recs = [b"", b"x" * 5, os.urandom(1000)]
    inputs.write_tfrecords(path, recs)
    assert list(inputs.read_tfrecords(path)) == recs
    raw = bytearray(open(path, "rb").read())
    assert struct.unpack("<Q", raw[:8])[0] == 0 and len(raw) == sum(16 + len(r) for r in recs)
    raw[-6] ^= 1
    open(path, "wb").write(raw)
    with pytest.raises(ValueError, match="corrupt"):
        list(inputs.read_tfrecords(path))
    assert len(list(inputs.read_tfrecords(path, verify_crc=False))) == 3
    open(path, "wb").write(raw[:-3])
    with pytest.raises(ValueError, match="truncated"):
        list(inputs.read_tfrecords(path, verify_crc=False))


def test_example_wire_format():
    # hand-assembled Example: features{feature{key:"text" value{bytes_list{value:"hi"}}}}
    hand = bytes([0x0A, 0x10, 0x0A, 0x0E, 0x0A, 0x04]) + b"text" + bytes([0x12, 0x06, 0x0A, 0x04, 0x0A, 0x02]) + b"hi"
    assert inputs.make_example("hi") == hand
    assert inputs.parse_example(hand) == {"text": [b"hi"]}
    ints = [0, 1, 127, 128, 300, 2 ** 40, -1]
    ex = inputs.make_example(ints)
    assert inputs.parse_example(ex)["text"].tolist() == ints
    # unpacked int64_list (one varint field per value) must parse as well
    lst = b"".join(bytes([0x08]) + inputs._put_varint(v) for v in (5, 6, 70000))
    feat = inputs._len_field(3, lst)
    entry = inputs._len_field(1, b"text") + inputs._len_field(2, feat)
    assert inputs.parse_example(inputs._len_field(1, inputs._len_field(1, entry)))["text"].tolist() == [5, 6, 70000]
    assert inputs.decode_bytestring(inputs.make_example("aé€😀")).tolist() == [0x61, 0xE9, 0x20AC, 0x1F600]
    assert inputs.decode_intstring(ex).dtype == np.int32


def test_windows_follow_the_reference_window_shift_rule():
    toks = np.arange(23, dtype=np.int32)
    w = inputs.text_windows(toks, ctx=5, patch_size=1)       # window(size=6, shift=5, drop_remainder=True)
    assert w.tolist() == [list(range(i, i + 6)) for i in (0, 5, 10, 15)]
    assert inputs.text_windows(toks[:5], 5, 1).shape == (0, 6)
    assert inputs.text_windows(toks, 5, 1, skip=3)[0].tolist() == list(range(3, 9))
    assert inputs.text_windows(toks[:6], 5, 1).tolist() == [list(range(6))]


def test_interleave_matches_tf_data_semantics():
    src = lambda name, n: (f"{name}{i}" for i in range(n))  # noqa: E731
    out = list(inputs.interleave([src("a", 3), src("b", 1), src("c", 2), src("d", 2)], 2))
    # slots (a, b): a0 b0 a1 [b exhausted -> c] c0 a2 c1 [a exhausted -> d] ... as Dataset.interleave(cycle_length=2)
    assert out == ["a0", "b0", "a1", "c0", "a2", "c1", "d0", "d1"]
    assert list(inputs.interleave([src("a", 2), src("b", 2)], 1)) == ["a0", "a1", "b0", "b1"]
    assert list(inputs.interleave([], 4)) == []


def test_split_files_shards_and_shuffles_like_the_reference():
    import random
    files = [f"f{i}" for i in range(7)]
    assert inputs.split_files(files[::-1], 1, 3, 0)[0] == ["f1", "f4"]
    expect = sorted(files)
    random.seed(99)
    random.shuffle(expect)
    assert inputs.split_files(files, 0, 2, 99)[0] == expect[0::2]
    with pytest.raises(ValueError):
        inputs.split_files([], 0, 1, 0)


def test_gpt_neo_input_end_to_end(tmp_path):
    cfg = mixer_config("tiny")
    cfg.update(sequence_length=8, shuffle_input_filenames=False, interleaved_datasets=2,
               dataset_configs=[{"path": str(tmp_path / "*int64*.tfrecord"), "type": "text", "weight": 1}])
    params = ModelParameter(cfg)
    docs_a = [list(range(100, 120)), list(range(200, 209))]           # windows: 2 (+ remainder dropped), 1
    docs_b = [list(range(300, 317))]                                  # windows: 2
    inputs.write_text_tfrecord(str(tmp_path / "a_int64_0.tfrecord"), docs_a)
    inputs.write_text_tfrecord(str(tmp_path / "b_int64_1.tfrecord"), docs_b)
    batches = list(inputs.gpt_neo_input(params, sub_batch_size=2, slice_index=0, slice_count=1))
    # interleave(cycle 2): a:w0 b:w0 a:w1 b:w1 a:w2 -> batches (a0,b0), (a1,b1); the odd last window is dropped
    assert len(batches) == 2
    x, y = batches[0]["token_x"], batches[0]["token_y"]
    assert tuple(x.shape) == (2, 8, 1) and x.dtype.is_floating_point is False
    assert x[:, :, 0].tolist() == [list(range(100, 108)), list(range(300, 308))]
    assert y[:, :, 0].tolist() == [list(range(101, 109)), list(range(301, 309))]        # output_offset = 1
    assert batches[1]["token_x"][:, 0, 0].tolist() == [108, 308]                         # windows advance by ctx
    # byte-string files decode code points; two workers see disjoint files
    cfg["dataset_configs"] = [{"path": str(tmp_path / "*text*.tfrecord"), "type": "text", "weight": 1}]
    params = ModelParameter(cfg)
    inputs.write_text_tfrecord(str(tmp_path / "text_0.tfrecord"), ["abcdefghijklmnopqr"])
    inputs.write_text_tfrecord(str(tmp_path / "text_1.tfrecord"), ["ABCDEFGHIJKLMNOPQR"])
    w0 = list(inputs.gpt_neo_input(params, 1, 0, 2))
    w1 = list(inputs.gpt_neo_input(params, 1, 1, 2))
    assert bytes(w0[0]["token_x"][0, :, 0].tolist()) == b"abcdefgh" and bytes(w1[0]["token_y"][0, :, 0].tolist()) == b"BCDEFGHI"
    assert len(w0) == 2 and len(w1) == 2
