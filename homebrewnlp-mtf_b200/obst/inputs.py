"""Text input pipeline: the language branch of src/inputs.py (split_files :15-30, _text_decoder :231-250,
decode_bytestring / decode_intstring :253-268, gpt_neo_input :528-568) without TensorFlow.

The reference reads TFRecord files of `tf.train.Example` protos with one feature `text` -- either a UTF-8 byte string
(one token per code point) or an int64 list (one token per entry, file names containing "int64") -- cuts EVERY
RECORD into windows of `sequence_length + token_patch_size * output_offset` tokens that advance by `sequence_length`
(remainder dropped), batches `sub_batch_size` windows, and hands the model `token_x = w[:, :S]`,
`token_y = w[:, offset:S+offset]` (`_memory_func`).  This module restates exactly that as a Python generator of
int32 torch tensors in the layout get_train_model takes ([batch, sequence, token_patch]); the file format is
parsed by hand (TFRecord framing with masked CRC-32C, protobuf wire format of Example), so no TensorFlow and no
generated protobuf classes are needed.  `write_text_tfrecord` produces files the reference's pipeline reads.

Deterministic path only: `use_random_dataloader` (shuffle buffer, repeat) raises; a `runs_log` resumes where the
logged runs stopped (`simulate_data_pipeline`)."""
import glob
import random
import struct
import typing

import numpy as np
import torch

from .dataclass import ModelParameter

# ---------------------------------------------------------------------------------------------------------
# CRC-32C (Castagnoli), masked as TFRecord does
# ---------------------------------------------------------------------------------------------------------
_CRC_TABLE = []


def _crc_table():
    if not _CRC_TABLE:
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            _CRC_TABLE.append(c)
    return _CRC_TABLE


def crc32c_py(data: bytes) -> int:
    """Bytewise table form: the statement the native routine (obst_crc32c) is tested against."""
    table = _crc_table()
    c = 0xFFFFFFFF
    for b in data:
        c = table[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def crc32c(data: bytes) -> int:
    from . import native
    return native.crc32c(data)


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------------------
# TFRecord framing: uint64 length, uint32 masked crc(length), payload, uint32 masked crc(payload)
# ---------------------------------------------------------------------------------------------------------
def read_tfrecords(path: str, verify_crc: bool = True) -> typing.Iterator[bytes]:
    with open(path, "rb") as f:
        while True:
            head = f.read(12)
            if not head:
                return
            if len(head) < 12:
                raise ValueError(f"{path}: truncated record header")
            length, len_crc = struct.unpack("<QI", head)
            if verify_crc and masked_crc32c(head[:8]) != len_crc:
                raise ValueError(f"{path}: corrupt record length")
            payload = f.read(length)
            tail = f.read(4)
            if len(payload) < length or len(tail) < 4:
                raise ValueError(f"{path}: truncated record")
            if verify_crc and masked_crc32c(payload) != struct.unpack("<I", tail)[0]:
                raise ValueError(f"{path}: corrupt record payload")
            yield payload


def write_tfrecords(path: str, records: typing.Iterable[bytes]):
    with open(path, "wb") as f:
        for rec in records:
            head = struct.pack("<Q", len(rec))
            f.write(head + struct.pack("<I", masked_crc32c(head)) + rec + struct.pack("<I", masked_crc32c(rec)))


# ---------------------------------------------------------------------------------------------------------
# protobuf wire format of tf.train.Example { Features features = 1; }  Features { map<string, Feature> feature = 1; }
# Feature { oneof kind { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3; } }
# ---------------------------------------------------------------------------------------------------------
def _varint(buf: bytes, pos: int) -> typing.Tuple[int, int]:
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _fields(buf: bytes) -> typing.Iterator[typing.Tuple[int, int, typing.Union[int, bytes]]]:
    pos = 0
    while pos < len(buf):
        key, pos = _varint(buf, pos)
        num, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = _varint(buf, pos)
        elif wire == 2:
            n, pos = _varint(buf, pos)
            val = buf[pos:pos + n]
            pos += n
        elif wire == 1:
            val = buf[pos:pos + 8]
            pos += 8
        elif wire == 5:
            val = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wire}")
        yield num, wire, val


def parse_example(proto: bytes) -> typing.Dict[str, typing.Union[typing.List[bytes], np.ndarray]]:
    """feature name -> list of byte strings (bytes_list) or int64 array (int64_list)."""
    out = {}
    for num, _, features in _fields(proto):
        if num != 1:
            continue
        for fnum, _, entry in _fields(features):
            if fnum != 1:
                continue
            name, feature = None, b""
            for enum, _, val in _fields(entry):
                if enum == 1:
                    name = val.decode("utf-8")
                elif enum == 2:
                    feature = val
            for knum, _, lst in _fields(feature):
                if knum == 1:  # BytesList
                    out[name] = [v for n, _, v in _fields(lst) if n == 1]
                elif knum == 3:  # Int64List: packed (one LEN field of varints) or one varint field per value
                    vals = []
                    for n, wire, v in _fields(lst):
                        if n != 1:
                            continue
                        if wire == 2:
                            p = 0
                            while p < len(v):
                                x, p = _varint(v, p)
                                vals.append(x)
                        else:
                            vals.append(v)
                    arr = np.array(vals, dtype=np.uint64).astype(np.int64)  # two's complement for negatives
                    out[name] = arr
    return out


def _put_varint(x: int) -> bytes:
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _len_field(num: int, payload: bytes) -> bytes:
    return _put_varint((num << 3) | 2) + _put_varint(len(payload)) + payload


def make_example(text: typing.Union[str, bytes, typing.Sequence[int]]) -> bytes:
    """An Example with the single feature `text`: bytes_list for str / bytes, packed int64_list for integers."""
    if isinstance(text, str):
        text = text.encode("utf-8")
    if isinstance(text, bytes):
        feature = _len_field(1, _len_field(1, text))
    else:
        feature = _len_field(3, _len_field(1, b"".join(_put_varint(int(v)) for v in text)))
    entry = _len_field(1, b"text") + _len_field(2, feature)
    return _len_field(1, _len_field(1, entry))


def write_text_tfrecord(path: str, documents: typing.Iterable[typing.Union[str, bytes, typing.Sequence[int]]]):
    write_tfrecords(path, (make_example(d) for d in documents))


# ---------------------------------------------------------------------------------------------------------
# decoders and windowing
# ---------------------------------------------------------------------------------------------------------
def decode_bytestring(proto: bytes) -> np.ndarray:
    """src/inputs.py:253-257: the UTF-8 text as unicode code points, one token each."""
    text = parse_example(proto)["text"][0]
    return np.frombuffer(text.decode("utf-8").encode("utf-32-le"), dtype="<u4").astype(np.int32)


def decode_intstring(proto: bytes) -> np.ndarray:
    """src/inputs.py:260-267: the int64 list cast to int32."""
    return np.asarray(parse_example(proto)["text"]).astype(np.int32)


def text_windows(tokens: np.ndarray, ctx: int, patch_size: int, skip: int = 0) -> np.ndarray:
    """_text_decoder.chunk, src/inputs.py:241-248: windows of ctx + patch_size tokens advancing by ctx, remainder
    dropped.  Returns [n_windows, ctx + patch_size]."""
    tokens = tokens[skip:]
    size = ctx + patch_size
    n = 0 if len(tokens) < size else (len(tokens) - size) // ctx + 1
    if n == 0:
        return np.zeros((0, size), dtype=np.int32)
    idx = np.arange(n)[:, None] * ctx + np.arange(size)[None, :]
    return tokens[idx]


def interleave(sources: typing.Iterable[typing.Iterator], cycle_length: int) -> typing.Iterator:
    """Deterministic tf.data interleave with block_length 1 (src/inputs.py:554-559): up to `cycle_length` sources are
    open at once and yield one element each in turn; an exhausted source's slot is refilled from the input."""
    sources = iter(sources)
    slots: typing.List[typing.Optional[typing.Iterator]] = []
    for src in sources:
        slots.append(src)
        if len(slots) == cycle_length:
            break
    exhausted_input = len(slots) < cycle_length
    i = 0
    while any(s is not None for s in slots):
        src = slots[i]
        if src is not None:
            try:
                yield next(src)
            except StopIteration:
                nxt = None
                if not exhausted_input:
                    nxt = next(sources, None)
                    exhausted_input = nxt is None
                slots[i] = nxt
                continue  # the refilled slot produces in this same turn
        i = (i + 1) % len(slots)


def _record_tokens(path) -> int:
    """The element count a file name carries (`..._<count>.tfrecord`), read the way src/inputs.py:34 reads it: the last
    `_` field with the CHARACTERS of '.tfrecord' stripped from both ends."""
    return int(str(path).split("_")[-1].strip(".tfrecord"))


def simulate_data_pipeline(runs_log: typing.Sequence[dict], file_list: typing.Sequence) -> typing.Tuple[list, list]:
    """src/inputs.py:33-128: replays earlier runs over the (already ordered) file list to find where a resumed run
    starts.  Returns (drop, consumed): per file, whether it is left out altogether and how many leading tokens of
    each of its records are skipped.

    One earlier run = {steps, grad_accumulation, batch_size, slice_count, ctx, interleave_size, token_patch_size}.
    Every slice of that run read `steps * grad_accumulation * (batch_size // slice_count)` windows of `ctx` tokens,
    round-robin over groups of `interleave_size` of its files.  The replay keeps the reference's arithmetic as it is,
    including: a group the budget does not reach is still visited (it takes nothing, :106 tests the run's budget, not
    what is left of it); a group the budget covers whole has its count SET to the usable length (:115), a group it
    covers in part has the taken tokens ADDED (:103); files are dropped only by whole groups, and those groups are
    formed over ALL files of the slice, dropped ones included (:118-126), while reading groups the live ones."""
    size = [_record_tokens(f) for f in file_list]
    n = len(size)
    drop = [False] * n
    consumed = [0] * n
    for run in runs_log:
        live = [i for i in range(n) if not drop[i]]
        left = {i: size[i] - consumed[i] for i in live}
        slices, ctx, patch = run["slice_count"], run["ctx"], run["token_patch_size"]
        group = run["interleave_size"]
        quota = run["steps"] * run["grad_accumulation"] * (run["batch_size"] // slices)
        for s in range(slices):
            mine = live[s::slices]
            budget = quota
            for start in range(0, len(mine), group):
                members = mine[start:start + group]
                usable = [left[i] - ((left[i] - patch) % ctx) - patch for i in members]
                windows = sum(usable) // ctx
                if windows <= budget:  # the whole group was read
                    budget -= windows
                    for i, u in zip(members, usable):
                        drop[i] = True
                        consumed[i] = u
                    continue
                rest = list(usable)
                turn = 0
                while budget > 0 and sum(rest) > 0:
                    while rest[turn] <= 0:
                        turn = (turn + 1) % len(rest)
                    rest[turn] -= ctx
                    budget -= 1
                    turn = (turn + 1) % len(rest)
                for i, u, r in zip(members, usable, rest):
                    if r <= 0:
                        drop[i] = True
                    consumed[i] += u - r
                if quota <= 0:
                    break
        for s in range(slices):
            mine = list(range(n))[s::slices]
            for start in range(0, len(mine), group):
                members = mine[start:start + group]
                gone = sum(drop[i] for i in members) == group
                for i in members:
                    drop[i] = gone
    return drop, consumed


def split_files(filenames: typing.Sequence[str], slice_index: int, slice_count: int, seed: int, runs_log=None):
    """src/inputs.py:15-30."""
    if not filenames:
        raise ValueError
    files = sorted(filenames)
    if seed != 0:
        random.seed(seed)
        random.shuffle(files)
    skips = [0] * len(files)
    if runs_log is not None:
        drop, skips = simulate_data_pipeline(runs_log, files)
        files = [f for f, d in zip(files, drop) if not d]
        skips = [k for k, d in zip(skips, drop) if not d]
    return files[slice_index::slice_count], skips[slice_index::slice_count]


def gpt_neo_input(params: ModelParameter, sub_batch_size: int, slice_index: int, slice_count: int, runs_log=None,
                  verify_crc: bool = True) -> typing.Iterator[typing.Dict[str, torch.Tensor]]:
    """src/inputs.py:528-568, deterministic path: yields {'token_x', 'token_y'} int32 [sub_batch, S / patch, patch]."""
    if params.use_random_dataloader:
        raise NotImplementedError("use_random_dataloader (repeat + shuffle buffer) is outside the hot path")
    filenames = []
    for cfg in params.dataset_configs:
        filenames.extend(glob.glob(cfg["path"]))
    filenames, skips = split_files(filenames, slice_index, slice_count,
                                   params.shuffle_input_filenames * params.data_seed, runs_log)
    decoder = decode_intstring if "int64" in filenames[0] else decode_bytestring
    patch = params.token_patch_size
    offset = params.output_offset
    seq = params.sequence_length
    width = seq + patch * offset

    def file_windows(path, skip):
        for proto in read_tfrecords(path, verify_crc):
            for w in text_windows(decoder(proto), seq, patch * offset, skip):
                yield w

    windows = interleave((file_windows(p, s) for p, s in zip(filenames, skips)), params.interleaved_datasets)
    batch = []
    for w in windows:
        batch.append(w)
        if len(batch) == sub_batch_size:  # dset.batch(sub_batch_size) keeps a final partial batch...
            yield _memory_func(np.stack(batch), seq, patch, offset)
            batch = []
    if batch and len(batch) == sub_batch_size:
        yield _memory_func(np.stack(batch), seq, patch, offset)
    # ...which _memory_func's fixed-shape reshape (src/inputs.py:543-545) would reject: it is dropped here
    del width


def _memory_func(x: np.ndarray, seq: int, patch: int, offset: int) -> typing.Dict[str, torch.Tensor]:
    """src/inputs.py:543-551."""
    x = x.reshape(x.shape[0], seq // patch + offset, patch).astype(np.int32)
    if offset > 0:
        vals1, vals2 = x[:, :seq], x[:, offset:seq + offset]
    else:
        vals1 = vals2 = x
    return {"token_x": torch.from_numpy(np.ascontiguousarray(vals1)),
            "token_y": torch.from_numpy(np.ascontiguousarray(vals2))}
