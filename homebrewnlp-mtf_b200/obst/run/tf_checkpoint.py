"""TensorBundle (TensorFlow V2 checkpoint) reader / writer without TensorFlow.

The reference checkpoints through `tf.train.Saver(sharded=True)` over the MTF master variables
(src/run/run.py:158-175, src/run/utils_run.py:18-29): `<prefix>.index` + `<prefix>.data-00000-of-00001` + the
`checkpoint` state file, every variable under its graph name (src/utils_mtf.py:271-280), every optimizer slot under
`<variable>/<optimizer string>/<slot>` (src/optimizer/backend.py:23-25).  TensorFlow is not available on this stack
(tensorflow 2.7 is a third-party dependency of the reference, requirements.txt), so the on-disk format is restated
here from its published definition:

  * `.data-*`: the tensors' raw little-endian bytes back to back;
  * `.index`: an immutable sorted string table in the LevelDB table format (tensorflow/core/lib/io/table_builder.cc,
    format.cc): prefix-compressed blocks of (key, value) with restart points, each block followed by a 1-byte
    compression tag and a masked CRC-32C, an (empty) meta-index block, an index block of block handles and a 48-byte
    footer ending in the magic number 0xdb4775248b80fb57;
  * key "" -> BundleHeaderProto {num_shards, endianness, version}; key <tensor name> -> BundleEntryProto {dtype, shape,
    shard_id, offset, size, masked crc32c of the bytes} (tensorflow/core/protobuf/tensor_bundle.proto).

PARITY UNPINNED: no TensorFlow-written file exists on this stack to read, nor TensorFlow to read what this writes;
the tests pin the pieces that have published known answers (CRC-32C check values, the table magic, protobuf wire
encodings) and the round trip.  Checksums go through the native obst_crc32c (gigabytes per save)."""
import os
import struct
import typing

import numpy as np
import torch

from .. import native

TABLE_MAGIC = 0xdb4775248b80fb57
BLOCK_SIZE = 262144          # table::Options::block_size
RESTART_INTERVAL = 16        # table::Options::block_restart_interval
HEADER_KEY = b""

# tensorflow/core/framework/types.proto
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_INT64, DT_BFLOAT16, DT_HALF = 1, 2, 3, 9, 14, 19
_TORCH_OF = {DT_FLOAT: torch.float32, DT_DOUBLE: torch.float64, DT_INT32: torch.int32, DT_INT64: torch.int64,
             DT_BFLOAT16: torch.bfloat16, DT_HALF: torch.float16}
_DT_OF = {v: k for k, v in _TORCH_OF.items()}


def mask_crc(crc: int) -> int:
    """crc32c::Mask: rotate right by 15 and add a constant (a CRC of data that embeds CRCs must not be a CRC)."""
    return ((((crc >> 15) | (crc << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def unmask_crc(masked: int) -> int:
    rot = (masked - 0xA282EAD8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------------------
# protobuf wire format (only what the two bundle messages need)
# ---------------------------------------------------------------------------------------------------------
def put_varint(x: int) -> bytes:
    if x < 0:
        x += 1 << 64  # int32 / int64 negative values are 10-byte two's-complement varints
    out = bytearray()
    while x >= 0x80:
        out.append((x & 0x7F) | 0x80)
        x >>= 7
    out.append(x)
    return bytes(out)


def get_varint(buf: bytes, pos: int) -> typing.Tuple[int, int]:
    x = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        x |= (b & 0x7F) << shift
        if b < 0x80:
            return x, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _tag(field: int, wire: int) -> bytes:
    return put_varint((field << 3) | wire)


def _fields(buf: bytes) -> typing.Iterator[typing.Tuple[int, int, typing.Union[int, bytes]]]:
    pos = 0
    while pos < len(buf):
        key, pos = get_varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = get_varint(buf, pos)
        elif wire == 1:
            val, pos = struct.unpack_from("<Q", buf, pos)[0], pos + 8
        elif wire == 2:
            n, pos = get_varint(buf, pos)
            val, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            val, pos = struct.unpack_from("<I", buf, pos)[0], pos + 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wire}")
        yield field, wire, val


def encode_header(num_shards: int = 1) -> bytes:
    """BundleHeaderProto{num_shards = 1, endianness = LITTLE (0: default, not written), version{producer = 1}}."""
    version = _tag(1, 0) + put_varint(1)  # VersionDef.producer = kTensorBundleVersion
    return _tag(1, 0) + put_varint(num_shards) + _tag(3, 2) + put_varint(len(version)) + version


def decode_header(buf: bytes) -> typing.Dict[str, int]:
    out = {"num_shards": 0, "endianness": 0, "producer": 0, "min_consumer": 0}
    for field, _, val in _fields(buf):
        if field == 1:
            out["num_shards"] = val
        elif field == 2:
            out["endianness"] = val
        elif field == 3:
            for f2, _, v2 in _fields(val):
                if f2 == 1:
                    out["producer"] = v2
                elif f2 == 2:
                    out["min_consumer"] = v2
    return out


def encode_entry(dtype: int, shape: typing.Sequence[int], shard_id: int, offset: int, size: int, crc_masked: int) -> bytes:
    """BundleEntryProto; proto3: zero-valued scalars are not written, a (possibly empty) shape message always is."""
    dims = b"".join(_tag(2, 2) + put_varint(len(d)) + d
                    for d in ((_tag(1, 0) + put_varint(int(n)) if n else b"") for n in shape))
    out = _tag(1, 0) + put_varint(dtype) + _tag(2, 2) + put_varint(len(dims)) + dims
    if shard_id:
        out += _tag(3, 0) + put_varint(shard_id)
    if offset:
        out += _tag(4, 0) + put_varint(offset)
    if size:
        out += _tag(5, 0) + put_varint(size)
    if crc_masked:
        out += _tag(6, 5) + struct.pack("<I", crc_masked)
    return out


def decode_entry(buf: bytes) -> typing.Dict[str, typing.Any]:
    out: typing.Dict[str, typing.Any] = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0, "crc32c": 0,
                                         "sliced": False}
    for field, _, val in _fields(buf):
        if field == 1:
            out["dtype"] = val
        elif field == 2:
            for f2, _, v2 in _fields(val):
                if f2 == 2:
                    size = 0
                    for f3, _, v3 in _fields(v2):
                        if f3 == 1:
                            size = v3 if v3 < 1 << 63 else v3 - (1 << 64)
                    out["shape"].append(size)
                elif f2 == 3 and v2:
                    raise ValueError("tensor of unknown rank in a checkpoint")
        elif field == 3:
            out["shard_id"] = val
        elif field == 4:
            out["offset"] = val
        elif field == 5:
            out["size"] = val
        elif field == 6:
            out["crc32c"] = val
        elif field == 7:
            out["sliced"] = True
    return out


# ---------------------------------------------------------------------------------------------------------
# the sorted string table
# ---------------------------------------------------------------------------------------------------------
class _BlockBuilder:
    def __init__(self, restart_interval: int):
        self.interval = restart_interval
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes):
        shared = 0
        if self.count < self.interval:
            limit = min(len(key), len(self.last_key))
            while shared < limit and key[shared] == self.last_key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.count = 0
        self.buf += put_varint(shared) + put_varint(len(key) - shared) + put_varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.count += 1

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self) -> bytes:
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))

    @property
    def empty(self) -> bool:
        return not self.buf


def _handle(offset: int, size: int) -> bytes:
    return put_varint(offset) + put_varint(size)


def build_table(items: typing.Sequence[typing.Tuple[bytes, bytes]], block_size: int = BLOCK_SIZE) -> bytes:
    """Serialise sorted (key, value) pairs as a table file (TableBuilder with no compression)."""
    keys = [k for k, _ in items]
    if keys != sorted(keys) or len(set(keys)) != len(keys):
        raise ValueError("table keys must be strictly increasing")
    out = bytearray()

    def write_block(contents: bytes) -> typing.Tuple[int, int]:
        offset = len(out)
        trailer_crc = mask_crc(native.crc32c(contents + b"\x00"))
        out.extend(contents)
        out.append(0)  # kNoCompression
        out.extend(struct.pack("<I", trailer_crc))
        return offset, len(contents)

    index = _BlockBuilder(1)
    block = _BlockBuilder(RESTART_INTERVAL)
    for key, value in items:
        block.add(key, value)
        if block.size_estimate() >= block_size:
            # the index key may be any string >= the block's last key and < the next block's first: the last key is
            index.add(block.last_key, _handle(*write_block(block.finish())))
            block = _BlockBuilder(RESTART_INTERVAL)
    if not block.empty:
        index.add(block.last_key, _handle(*write_block(block.finish())))
    meta = _handle(*write_block(_BlockBuilder(RESTART_INTERVAL).finish()))
    idx = _handle(*write_block(index.finish()))
    footer = meta + idx
    footer += bytes(40 - len(footer))
    footer += struct.pack("<II", TABLE_MAGIC & 0xFFFFFFFF, TABLE_MAGIC >> 32)
    out.extend(footer)
    return bytes(out)


def _read_block(data: bytes, offset: int, size: int, verify: bool) -> typing.List[typing.Tuple[bytes, bytes]]:
    contents = data[offset:offset + size]
    if len(contents) != size or offset + size + 5 > len(data):
        raise ValueError("truncated table block")
    ctype = data[offset + size]
    if ctype != 0:
        raise NotImplementedError("compressed table blocks (TensorBundle writes none)")
    if verify:
        stored = struct.unpack_from("<I", data, offset + size + 1)[0]
        if unmask_crc(stored) != native.crc32c(data[offset:offset + size + 1]):
            raise ValueError(f"table block at {offset}: checksum mismatch")
    n_restarts = struct.unpack_from("<I", contents, size - 4)[0]
    end = size - 4 - 4 * n_restarts
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = get_varint(contents, pos)
        unshared, pos = get_varint(contents, pos)
        vlen, pos = get_varint(contents, pos)
        key = key[:shared] + contents[pos:pos + unshared]
        pos += unshared
        out.append((key, contents[pos:pos + vlen]))
        pos += vlen
    return out


def read_table(data: bytes, verify: bool = True) -> typing.List[typing.Tuple[bytes, bytes]]:
    if len(data) < 48:
        raise ValueError("not a table file: shorter than its footer")
    footer = data[-48:]
    lo, hi = struct.unpack("<II", footer[40:])
    if (hi << 32) | lo != TABLE_MAGIC:
        raise ValueError("not a table file: bad magic number")
    pos = 0
    _, pos = get_varint(footer, pos)  # meta-index handle (unused)
    _, pos = get_varint(footer, pos)
    ioff, pos = get_varint(footer, pos)
    isize, pos = get_varint(footer, pos)
    out: typing.List[typing.Tuple[bytes, bytes]] = []
    for _, handle in _read_block(data, ioff, isize, verify):
        boff, p = get_varint(handle, 0)
        bsize, _ = get_varint(handle, p)
        out.extend(_read_block(data, boff, bsize, verify))
    return out


# ---------------------------------------------------------------------------------------------------------
# bundles
# ---------------------------------------------------------------------------------------------------------
def _host_bytes(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous().reshape(-1)
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint8)
    return t.numpy().view(np.uint8)


def write_bundle(prefix: str, tensors: typing.Dict[str, torch.Tensor]):
    """`prefix`.index + `prefix`.data-00000-of-00001 holding `tensors` (name -> tensor, any device)."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    entries = []
    offset = 0
    tmp = f"{prefix}.data-00000-of-00001.tmp"
    with open(tmp, "wb") as f:
        for name in sorted(tensors):  # data in key order: one sequential read restores everything
            t = tensors[name]
            if t.dtype not in _DT_OF:
                raise TypeError(f"checkpoint tensor '{name}' has unsupported dtype {t.dtype}")
            raw = _host_bytes(t)
            crc = mask_crc(native.crc32c(raw)) if raw.size else mask_crc(0)
            f.write(memoryview(raw))
            key = name.encode("utf-8")
            if not key:
                raise ValueError("the empty name is the bundle header's key")
            entries.append((key, encode_entry(_DT_OF[t.dtype], list(t.shape), 0, offset, raw.size, crc)))
            offset += raw.size
    os.replace(tmp, f"{prefix}.data-00000-of-00001")
    table = build_table([(HEADER_KEY, encode_header(1))] + sorted(entries))
    with open(prefix + ".index.tmp", "wb") as f:
        f.write(table)
    os.replace(prefix + ".index.tmp", prefix + ".index")


def read_bundle(prefix: str, verify: bool = True, names: typing.Optional[typing.Container[str]] = None
                ) -> typing.Dict[str, torch.Tensor]:
    with open(prefix + ".index", "rb") as f:
        items = read_table(f.read(), verify)
    if not items or items[0][0] != HEADER_KEY:
        raise ValueError("bundle index has no header entry")
    header = decode_header(items[0][1])
    if header["endianness"] != 0:
        raise NotImplementedError("big-endian bundle")
    if header["min_consumer"] > 1:
        raise NotImplementedError(f"bundle needs a reader of version {header['min_consumer']}")
    shards: typing.Dict[int, typing.Any] = {}
    out: typing.Dict[str, torch.Tensor] = {}
    try:
        for key, value in items[1:]:
            name = key.decode("utf-8")
            if names is not None and name not in names:
                continue
            e = decode_entry(value)
            if e["sliced"]:
                raise NotImplementedError(f"'{name}' is stored as partitioned slices")
            if e["dtype"] not in _TORCH_OF:
                raise TypeError(f"'{name}': unsupported DataType {e['dtype']}")
            if e["shard_id"] not in shards:
                path = f"{prefix}.data-{e['shard_id']:05d}-of-{header['num_shards']:05d}"
                shards[e["shard_id"]] = open(path, "rb")
            f = shards[e["shard_id"]]
            f.seek(e["offset"])
            raw = np.frombuffer(f.read(e["size"]), dtype=np.uint8)
            if raw.size != e["size"]:
                raise ValueError(f"'{name}': data file is truncated")
            if verify and unmask_crc(e["crc32c"]) != (native.crc32c(raw) if raw.size else 0):
                raise ValueError(f"'{name}': checksum mismatch")
            dt = _TORCH_OF[e["dtype"]]
            count = int(np.prod(e["shape"])) if e["shape"] else 1
            if count * torch.empty((), dtype=dt).element_size() != raw.size:
                raise ValueError(f"'{name}': {raw.size} bytes do not fill shape {e['shape']} of {dt}")
            if raw.size == 0:
                t = torch.empty(0, dtype=dt)
            elif dt == torch.bfloat16:
                t = torch.from_numpy(raw.view(np.int16).copy()).view(torch.bfloat16)
            else:
                t = torch.from_numpy(raw.copy()).view(dt)
            out[name] = t.reshape(e["shape"])
    finally:
        for f in shards.values():
            f.close()
    return out


def write_checkpoint_state(directory: str, latest: str, others: typing.Sequence[str] = ()):
    """The `checkpoint` text-proto CheckpointState file tf.train.latest_checkpoint reads."""
    paths = list(others) + ([latest] if latest not in others else [])
    with open(os.path.join(directory, "checkpoint"), "w") as f:
        f.write(f'model_checkpoint_path: "{latest}"\n')
        for p in paths:
            f.write(f'all_model_checkpoint_paths: "{p}"\n')


def latest_checkpoint(directory: str) -> typing.Optional[str]:
    path = os.path.join(directory, "checkpoint")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        for line in f:
            if line.startswith("model_checkpoint_path:"):
                name = line.split(":", 1)[1].strip().strip('"')
                return name if os.path.isabs(name) else os.path.join(directory, name)
    return None
