"""The hand-written file formats (TFRecord framing + CRC-32C, `tf.train.Example`, the TensorBundle protos of
obst/run/tf_checkpoint.py) against INDEPENDENT implementations that happen to be in this image:

  * TensorBoard's TensorFlow stub (`tensorboard.compat.tensorflow_stub.pywrap_tensorflow`: `crc32c`, `masked_crc32c`,
    `PyRecordReader_New`) and its `RecordWriter` -- the TensorFlow team's own pure-Python TFRecord reader / writer;
  * the protobuf runtime (`google.protobuf`) with TensorFlow's compiled `TensorShapeProto` / `DataType` / `VersionDef`
    (`tensorboard.compat.proto`, package renamed `tensorboard`, same wire format) and `BundleHeaderProto` /
    `BundleEntryProto` / `Example` declared here from the published tensor_bundle.proto / example.proto / feature.proto.

Neither TensorFlow nor a TensorFlow-written checkpoint exists on this stack, so the LevelDB-format table of the `.index`
file stays pinned only by its published constants (tests/test_tf_checkpoint.py); everything INSIDE it and the whole
input-file format is cross-checked here.  CPU only."""
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))

pw = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
record_writer = pytest.importorskip("tensorboard.summary.writer.record_writer")
descriptor_pb2 = pytest.importorskip("google.protobuf.descriptor_pb2")


def _message_classes():
    from google.protobuf import descriptor_pool, message_factory
    from tensorboard.compat.proto import tensor_shape_pb2, types_pb2, versions_pb2
    pool = descriptor_pool.DescriptorPool()
    for mod in (tensor_shape_pb2, types_pb2, versions_pb2):
        pool.AddSerializedFile(mod.DESCRIPTOR.serialized_pb)
    F = descriptor_pb2.FieldDescriptorProto

    def field(msg, name, number, ftype, label=F.LABEL_OPTIONAL, type_name=None, packed=None):
        f = msg.field.add(name=name, number=number, type=ftype, label=label)
        if type_name:
            f.type_name = type_name
        if packed is not None:
            f.options.packed = packed
        return f

    fd = descriptor_pb2.FileDescriptorProto(name="obst_test/tensor_bundle.proto", package="obst_test", syntax="proto3")
    fd.dependency.extend([tensor_shape_pb2.DESCRIPTOR.name, types_pb2.DESCRIPTOR.name, versions_pb2.DESCRIPTOR.name])
    header = fd.message_type.add(name="BundleHeaderProto")  # tensorflow/core/protobuf/tensor_bundle.proto
    endian = header.enum_type.add(name="Endianness")
    endian.value.add(name="LITTLE", number=0)
    endian.value.add(name="BIG", number=1)
    field(header, "num_shards", 1, F.TYPE_INT32)
    field(header, "endianness", 2, F.TYPE_ENUM, type_name=".obst_test.BundleHeaderProto.Endianness")
    field(header, "version", 3, F.TYPE_MESSAGE, type_name=".tensorboard.VersionDef")
    entry = fd.message_type.add(name="BundleEntryProto")
    field(entry, "dtype", 1, F.TYPE_ENUM, type_name=".tensorboard.DataType")
    field(entry, "shape", 2, F.TYPE_MESSAGE, type_name=".tensorboard.TensorShapeProto")
    field(entry, "shard_id", 3, F.TYPE_INT32)
    field(entry, "offset", 4, F.TYPE_INT64)
    field(entry, "size", 5, F.TYPE_INT64)
    field(entry, "crc32c", 6, F.TYPE_FIXED32)
    # tensorflow/core/example/feature.proto, example.proto
    bytes_list = fd.message_type.add(name="BytesList")
    field(bytes_list, "value", 1, F.TYPE_BYTES, F.LABEL_REPEATED)
    float_list = fd.message_type.add(name="FloatList")
    field(float_list, "value", 1, F.TYPE_FLOAT, F.LABEL_REPEATED, packed=True)
    int64_list = fd.message_type.add(name="Int64List")
    field(int64_list, "value", 1, F.TYPE_INT64, F.LABEL_REPEATED, packed=True)
    feature = fd.message_type.add(name="Feature")
    feature.oneof_decl.add(name="kind")
    for i, (name, tname) in enumerate((("bytes_list", "BytesList"), ("float_list", "FloatList"),
                                       ("int64_list", "Int64List"))):
        field(feature, name, i + 1, F.TYPE_MESSAGE, type_name=f".obst_test.{tname}").oneof_index = 0
    features = fd.message_type.add(name="Features")
    map_entry = features.nested_type.add(name="FeatureEntry")
    map_entry.options.map_entry = True
    field(map_entry, "key", 1, F.TYPE_STRING)
    field(map_entry, "value", 2, F.TYPE_MESSAGE, type_name=".obst_test.Feature")
    field(features, "feature", 1, F.TYPE_MESSAGE, F.LABEL_REPEATED, type_name=".obst_test.Features.FeatureEntry")
    example = fd.message_type.add(name="Example")
    field(example, "features", 1, F.TYPE_MESSAGE, type_name=".obst_test.Features")
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("obst_test." + n))  # noqa: E731
    return get("BundleHeaderProto"), get("BundleEntryProto"), get("Example"), types_pb2


def test_crc32c_matches_tensorboards_implementation():
    from obst import native
    from obst.inputs import crc32c, crc32c_py, masked_crc32c
    from obst.run.tf_checkpoint import mask_crc, unmask_crc
    rng = np.random.default_rng(3)
    for n in [0, 1, 2, 3, 7, 8, 9, 15, 16, 17, 63, 64, 65, 255, 1000, 4096, 12289, 100003]:
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        want = pw.crc32c(data)
        assert crc32c(data) == want == crc32c_py(data), n
        assert masked_crc32c(data) == pw.masked_crc32c(data) == mask_crc(want), n
        assert unmask_crc(pw.masked_crc32c(data)) == want, n
        # the running form the checkpoint writer uses (a tensor checksummed in pieces)
        cut = n // 3
        assert native.crc32c(data[cut:], native.crc32c(data[:cut])) == want, n


def test_tfrecord_files_are_exchangeable_with_tensorboards_reader_and_writer(tmp_path):
    from obst.inputs import read_tfrecords, write_tfrecords
    rng = np.random.default_rng(4)
    records = [rng.integers(0, 256, int(n), dtype=np.uint8).tobytes() for n in [0, 1, 5, 100, 4096, 70000]]
    mine = str(tmp_path / "mine.tfrecord")
    write_tfrecords(mine, records)
    reader = pw.PyRecordReader_New(mine)
    got = []
    while True:
        try:
            reader.GetNext()
        except pw.errors.OutOfRangeError:
            break
        got.append(bytes(reader.record()))
    assert got == records
    theirs = str(tmp_path / "theirs.tfrecord")
    with open(theirs, "wb") as f:
        w = record_writer.RecordWriter(f)
        for r in records:
            w.write(r)
        w.flush()
    assert list(read_tfrecords(theirs, verify_crc=True)) == records
    assert open(theirs, "rb").read() == open(mine, "rb").read()
    # a flipped payload bit is caught by both
    blob = bytearray(open(mine, "rb").read())
    blob[16 + 17 + 12 + 2] ^= 0x10  # records of 0 and 1 bytes (12 + n + 4 each), the third one's header, its third byte
    bad = str(tmp_path / "bad.tfrecord")
    open(bad, "wb").write(bytes(blob))
    with pytest.raises(Exception):
        list(read_tfrecords(bad, verify_crc=True))


def test_example_protos_match_the_protobuf_runtime():
    from obst.inputs import decode_bytestring, decode_intstring, make_example, parse_example
    _, _, Example, _ = _message_classes()
    rng = np.random.default_rng(5)
    # mine -> protobuf
    text = "naïve café ☕ text\n".encode("utf-8")
    msg = Example.FromString(make_example(text))
    assert list(msg.features.feature) == ["text"] and list(msg.features.feature["text"].bytes_list.value) == [text]
    tokens = [int(v) for v in rng.integers(0, 1 << 40, 300)] + [0, 1, 127, 128, (1 << 63) - 1]
    msg = Example.FromString(make_example(tokens))
    assert list(msg.features.feature["text"].int64_list.value) == tokens
    # protobuf -> mine (packed, as every current writer emits them, and negative values)
    msg = Example()
    msg.features.feature["text"].int64_list.value.extend(tokens + [-1, -(1 << 40)])
    msg.features.feature["other"].float_list.value.extend([1.0, 2.5])
    blob = msg.SerializeToString()
    assert parse_example(blob)["text"].tolist() == tokens + [-1, -(1 << 40)]
    assert decode_intstring(blob).tolist() == np.asarray(tokens + [-1, -(1 << 40)]).astype(np.int32).tolist()
    msg = Example()
    msg.features.feature["text"].bytes_list.value.append(text)
    assert decode_bytestring(msg.SerializeToString()).tolist() == [ord(c) for c in text.decode("utf-8")]
    # and byte for byte where the encoding is canonical (one feature: no map-order freedom)
    assert msg.SerializeToString(deterministic=True) == make_example(text)


def test_tensor_bundle_protos_match_the_protobuf_runtime():
    from obst.run import tf_checkpoint as tfc
    Header, Entry, _, types_pb2 = _message_classes()
    # dtype codes are TensorFlow's enum values
    assert (tfc.DT_FLOAT, tfc.DT_DOUBLE, tfc.DT_INT32, tfc.DT_INT64, tfc.DT_BFLOAT16) == (
        types_pb2.DT_FLOAT, types_pb2.DT_DOUBLE, types_pb2.DT_INT32, types_pb2.DT_INT64, types_pb2.DT_BFLOAT16)
    # header: mine -> protobuf, byte-identical to what protobuf writes for the same message
    h = Header.FromString(tfc.encode_header(1))
    assert (h.num_shards, h.endianness, h.version.producer, h.version.min_consumer) == (1, 0, 1, 0)
    ref = Header(num_shards=1)
    ref.version.producer = 1
    assert ref.SerializeToString(deterministic=True) == tfc.encode_header(1)
    assert tfc.decode_header(ref.SerializeToString()) == {"num_shards": 1, "endianness": 0, "producer": 1,
                                                          "min_consumer": 0}
    # entries, both directions, over shapes with zeros and large offsets
    rng = np.random.default_rng(6)
    cases = [(tfc.DT_FLOAT, [], 0, 0, 4, 0x12345678), (tfc.DT_BFLOAT16, [8, 0, 3], 0, 1 << 33, 0, 0),
             (tfc.DT_INT64, [1], 0, 7, 8, 0xFFFFFFFF)]
    for _ in range(20):
        shape = [int(v) for v in rng.integers(0, 5000, int(rng.integers(0, 5)))]
        cases.append((int(rng.choice([tfc.DT_FLOAT, tfc.DT_BFLOAT16, tfc.DT_INT32])), shape, 0,
                      int(rng.integers(0, 1 << 40)), int(rng.integers(0, 1 << 34)), int(rng.integers(0, 1 << 32))))
    for dtype, shape, shard, offset, size, crc in cases:
        blob = tfc.encode_entry(dtype, shape, shard, offset, size, crc)
        e = Entry.FromString(blob)
        assert (e.dtype, [d.size for d in e.shape.dim], e.shard_id, e.offset, e.size, e.crc32c) == (
            dtype, shape, shard, offset, size, crc)
        ref = Entry(dtype=dtype, shard_id=shard, offset=offset, size=size, crc32c=crc)
        ref.shape.SetInParent()
        for n in shape:
            ref.shape.dim.add(size=n)
        assert ref.SerializeToString(deterministic=True) == blob, (dtype, shape, offset, size, crc)
        d = tfc.decode_entry(ref.SerializeToString())
        assert (d["dtype"], d["shape"], d["shard_id"], d["offset"], d["size"], d["crc32c"], d["sliced"]) == (
            dtype, shape, shard, offset, size, crc, False)


def test_a_written_bundle_holds_protobuf_readable_entries(tmp_path):
    """End to end: every value of the `.index` table of a bundle written by write_bundle parses with the protobuf
    runtime, and (offset, size, masked crc) describe the `.data` file -- checked with TensorBoard's CRC."""
    import torch
    from obst.run import tf_checkpoint as tfc
    Header, Entry, _, _ = _message_classes()
    g = torch.Generator().manual_seed(0)
    tensors = {"a/b": torch.randn(3, 5, generator=g), "a/c": torch.randn(7, generator=g).to(torch.bfloat16),
               "step": torch.tensor(12345, dtype=torch.int64), "empty": torch.zeros(0, 4)}
    prefix = str(tmp_path / "model.ckpt-1")
    tfc.write_bundle(prefix, tensors)
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    items = dict(tfc.read_table(open(prefix + ".index", "rb").read()))
    h = Header.FromString(items.pop(b""))
    assert h.num_shards == 1 and h.version.producer == 1
    assert sorted(items) == sorted(k.encode() for k in tensors)
    for key, blob in items.items():
        e = Entry.FromString(blob)
        t = tensors[key.decode()]
        assert [d.size for d in e.shape.dim] == list(t.shape)
        assert e.size == t.numel() * t.element_size()
        assert pw.masked_crc32c(data[e.offset:e.offset + e.size]) == e.crc32c
        raw = t.reshape(-1).contiguous().view(torch.uint8).numpy().tobytes() if t.numel() else b""
        assert data[e.offset:e.offset + e.size] == raw
    assert struct.unpack("<Q", open(prefix + ".index", "rb").read()[-8:])[0] == tfc.TABLE_MAGIC
