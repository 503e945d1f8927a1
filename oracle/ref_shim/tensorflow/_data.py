"""tf.data / TFRecord / tf.Example stand-ins (see ../README.md): datasets are re-iterable Python generators with the
deterministic semantics of tf.data's sequential transformations -- `interleave` opens `cycle_length` inputs and takes
`block_length` elements from each in turn, replacing an exhausted one by the next input; `window(size, shift, stride,
drop_remainder)` yields datasets of `size` consecutive elements starting every `shift`; `batch` stacks."""
import glob as _glob
import struct

import numpy as np


class Dataset:
    def __init__(self, make_iter):
        self._make_iter = make_iter

    def __iter__(self):
        return self._make_iter()

    # -- sources -------------------------------------------------------------------------------------------
    @staticmethod
    def from_tensor_slices(x):
        if isinstance(x, tuple):
            return Dataset.zip(tuple(Dataset.from_tensor_slices(e) for e in x))
        items = list(x) if not isinstance(x, np.ndarray) else [x[i] for i in range(x.shape[0])]
        return Dataset(lambda: iter(items))

    @staticmethod
    def from_tensors(x):
        return Dataset(lambda: iter([x]))

    @staticmethod
    def zip(datasets):
        datasets = tuple(datasets)
        return Dataset(lambda: iter(zip(*[iter(d) for d in datasets])))

    # -- transformations -----------------------------------------------------------------------------------
    def skip(self, count):
        count = int(np.asarray(count))

        def gen():
            it = iter(self)
            for _ in range(count):
                if next(it, _END) is _END:
                    return
            yield from it
        return Dataset(gen)

    def take(self, count):
        def gen():
            for i, x in enumerate(self):
                if i >= count:
                    return
                yield x
        return Dataset(gen)

    def repeat(self, count=None):
        def gen():
            n = 0
            while count is None or n < count:
                yield from self
                n += 1
        return Dataset(gen)

    def map(self, fn, num_parallel_calls=None, deterministic=None):
        return Dataset(lambda: (_call(fn, x) for x in self))

    def batch(self, batch_size, drop_remainder=False, **kwargs):
        batch_size = int(batch_size)

        def gen():
            buf = []
            for x in self:
                buf.append(x)
                if len(buf) == batch_size:
                    yield _stack(buf)
                    buf = []
            if buf and not drop_remainder:
                yield _stack(buf)
        return Dataset(gen)

    def window(self, size, shift=None, stride=1, drop_remainder=False):
        shift = size if shift is None else shift

        def gen():
            items = list(self)  # sequential semantics; the documents here are short
            start = 0
            while start < len(items):
                win = items[start:start + size * stride:stride]
                if len(win) < size and drop_remainder:
                    return
                yield Dataset(lambda win=win: iter(win))
                start += shift
        return Dataset(gen)

    def interleave(self, map_func, cycle_length=1, block_length=1, num_parallel_calls=None, deterministic=None):
        def gen():
            inputs = iter(self)
            slots = []
            exhausted_inputs = False

            def refill():
                nonlocal exhausted_inputs
                while len(slots) < cycle_length and not exhausted_inputs:
                    x = next(inputs, _END)
                    if x is _END:
                        exhausted_inputs = True
                    else:
                        slots.append(iter(_call(map_func, x)))
            refill()
            i = 0
            while slots:
                i %= len(slots)
                it = slots[i]
                done = False
                for _ in range(block_length):
                    x = next(it, _END)
                    if x is _END:
                        done = True
                        break
                    yield x
                if done:
                    # the exhausted slot is replaced IN PLACE by the next input; with none left the cycle shrinks
                    nxt = _END if exhausted_inputs else next(inputs, _END)
                    if nxt is _END:
                        exhausted_inputs = True
                        slots.pop(i)
                    else:
                        slots[i] = iter(_call(map_func, nxt))
                else:
                    i += 1
        return Dataset(gen)

    def shuffle(self, *a, **k):
        raise NotImplementedError("shuffle buffers are not reproduced by the stand-in")

    def prefetch(self, *a, **k):
        return self


_END = object()


def _call(fn, x):
    return fn(*x) if isinstance(x, tuple) else fn(x)


def _stack(items):
    if isinstance(items[0], tuple):
        return tuple(_stack([it[k] for it in items]) for k in range(len(items[0])))
    if isinstance(items[0], dict):
        return {k: _stack([it[k] for it in items]) for k in items[0]}
    return np.stack([np.asarray(i) for i in items])


class TFRecordDataset(Dataset):
    """uint64 length, uint32 masked crc, payload, uint32 masked crc (checksums not verified here)."""

    def __init__(self, filenames, **kwargs):
        if isinstance(filenames, (str, bytes, np.str_)):
            filenames = [filenames]
        names = [f.decode() if isinstance(f, bytes) else str(f) for f in np.asarray(filenames).reshape(-1).tolist()]

        def gen():
            for name in names:
                with open(name, "rb") as f:
                    while True:
                        head = f.read(12)
                        if len(head) < 12:
                            break
                        (length,) = struct.unpack("<Q", head[:8])
                        payload = f.read(length)
                        f.read(4)
                        yield payload
        super().__init__(gen)


# ---- tf.Example --------------------------------------------------------------------------------------------
class FixedLenFeature:
    def __init__(self, shape, dtype, default_value=None):
        self.shape, self.dtype = shape, dtype


class VarLenFeature:
    def __init__(self, dtype):
        self.dtype = dtype


def _varint(buf, pos):
    x = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        x |= (b & 0x7F) << shift
        if b < 0x80:
            return x, pos
        shift += 7


def _fields(buf):
    pos = 0
    while pos < len(buf):
        key, pos = _varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = _varint(buf, pos)
        elif wire == 2:
            n, pos = _varint(buf, pos)
            val, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        elif wire == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        else:
            raise ValueError(f"wire type {wire}")
        yield field, wire, val


def parse_single_example(serialized, features):
    """Example{features{feature{key -> Feature{bytes_list | float_list | int64_list}}}}"""
    serialized = bytes(serialized)
    found = {}
    for f1, _, feats in _fields(serialized):
        if f1 != 1:
            continue
        for f2, _, entry in _fields(feats):
            if f2 != 1:
                continue
            key = value = None
            for f3, _, v in _fields(entry):
                if f3 == 1:
                    key = v.decode()
                elif f3 == 2:
                    value = v
            found[key] = value
    out = {}
    for name, spec in features.items():
        feature = found[name]
        result = None
        for kind, _, lst in _fields(feature):
            if kind == 1:  # BytesList
                vals = [v for f, _, v in _fields(lst) if f == 1]
                result = vals[0] if isinstance(spec, FixedLenFeature) and spec.shape == [] else vals
            elif kind == 3:  # Int64List (packed or not)
                vals = []
                for f, wire, v in _fields(lst):
                    if f != 1:
                        continue
                    if wire == 2:
                        pos = 0
                        while pos < len(v):
                            x, pos = _varint(v, pos)
                            vals.append(x if x < 1 << 63 else x - (1 << 64))
                    else:
                        vals.append(v if v < 1 << 63 else v - (1 << 64))
                result = np.asarray(vals, dtype=np.int64)
        out[name] = result
    return out


def unicode_decode(text, encoding):
    return np.asarray([ord(c) for c in bytes(text).decode(encoding)], dtype=np.int32)


def glob(pattern):
    return sorted(_glob.glob(pattern))
