"""NumPy stand-in for the slice of TensorFlow the reference's Python touches (see ../README.md).  Tensors are NumPy
arrays; float32 arithmetic stays float32 (NumPy >= 2 treats Python scalars as weak)."""
import builtins
import contextlib
import types

import numpy as np


class DType:
    def __init__(self, name, np_dtype, floating, bf16=False):
        self.name = name
        self.as_numpy_dtype = np_dtype
        self.is_floating = floating
        self.is_bf16 = bf16
        self.size = 2 if bf16 else np.dtype(np_dtype).itemsize

    def __repr__(self):
        return f"tf.{self.name}"

    @property
    def is_integer(self):
        return not self.is_floating and self.name != "bool"


float32 = DType("float32", np.float32, True)
float64 = DType("float64", np.float64, True)
float16 = DType("float16", np.float16, True)
bfloat16 = DType("bfloat16", np.float32, True, bf16=True)  # held as float32 values rounded to 8 significand bits
int32 = DType("int32", np.int32, False)
int64 = DType("int64", np.int64, False)
uint8 = DType("uint8", np.uint8, False)
bool = DType("bool", np.bool_, False)  # noqa: A001
dtypes = types.SimpleNamespace(DType=DType, float32=float32, bfloat16=bfloat16, int32=int32, int64=int64)


class Tensor:  # annotations only: values are numpy arrays
    pass


class Variable:
    pass


class Operation:
    pass


def round_bf16(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    bits = x.view(np.uint32)
    rounded = ((bits + 0x7FFF + ((bits >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
    out = rounded.view(np.float32).copy()
    out[np.isnan(x)] = np.nan
    return out.reshape(x.shape)


class _StaticShape(tuple):
    def as_list(self):
        return list(self)


class TFArray(np.ndarray):
    """ndarray whose `.shape` also answers `.as_list()`, as a tf.Tensor's static shape does."""

    @property
    def shape(self):
        return _StaticShape(np.ndarray.shape.__get__(self))

    # a tf.Tensor is immutable: `x += y` rebinds x to a new (broadcast) tensor
    def __iadd__(self, other): return self + other  # noqa: E704
    def __isub__(self, other): return self - other  # noqa: E704
    def __imul__(self, other): return self * other  # noqa: E704
    def __itruediv__(self, other): return self / other  # noqa: E704


def _w(x):
    return np.asarray(x).view(TFArray)


def cast(x, dtype, name=None):
    x = np.asarray(x)
    if not isinstance(dtype, DType):  # a NumPy dtype taken from another stand-in tensor
        return _w(x.astype(dtype))
    if dtype.is_bf16:
        return _w(round_bf16(x.astype(np.float32)))
    return _w(x.astype(dtype.as_numpy_dtype))


def tensor_scatter_nd_add(tensor, indices, updates, name=None):
    out = np.array(tensor, copy=True)
    idx = np.asarray(indices)
    np.add.at(out, tuple(idx[..., k] for k in builtins.range(idx.shape[-1])), np.asarray(updates))
    return _w(out)


def constant(value, dtype=None, shape=None, name=None):
    dtype = dtype or float32
    arr = cast(np.asarray(value), dtype)
    if shape is not None:
        arr = np.broadcast_to(arr, tuple(shape)).copy()
    return arr


def _f(fn):
    return lambda x, *a, name=None, **k: fn(np.asarray(x), *a, **k)


maximum = lambda a, b, name=None: np.maximum(a, b)  # noqa: E731
minimum = lambda a, b, name=None: np.minimum(a, b)  # noqa: E731
equal = lambda a, b, name=None: np.equal(a, b)  # noqa: E731
greater = lambda a, b, name=None: np.greater(a, b)  # noqa: E731
greater_equal = lambda a, b, name=None: np.greater_equal(a, b)  # noqa: E731
less = lambda a, b, name=None: np.less(a, b)  # noqa: E731
less_equal = lambda a, b, name=None: np.less_equal(a, b)  # noqa: E731
identity = lambda x, name=None: np.asarray(x)  # noqa: E731
stop_gradient = identity
reshape = lambda x, shape, name=None: _w(np.reshape(np.asarray(x), shape))  # noqa: E731
transpose = lambda x, perm=None, name=None: np.transpose(x, perm)  # noqa: E731
zeros = lambda shape, dtype=float32, name=None: np.zeros(shape, dtype.as_numpy_dtype)  # noqa: E731
ones = lambda shape, dtype=float32, name=None: np.ones(shape, dtype.as_numpy_dtype)  # noqa: E731
zeros_like = lambda x, name=None: np.zeros_like(x)  # noqa: E731
reduce_sum = lambda x, axis=None, keepdims=False, name=None: np.sum(x, axis=axis, keepdims=keepdims)  # noqa: E731
reduce_mean = lambda x, axis=None, keepdims=False, name=None: np.mean(x, axis=axis, keepdims=keepdims)  # noqa: E731
reduce_max = lambda x, axis=None, keepdims=False, name=None: np.max(x, axis=axis, keepdims=keepdims)  # noqa: E731
einsum = lambda eq, *xs, name=None: _w(np.einsum(eq, *[np.asarray(x) for x in xs]))  # noqa: E731
gather = lambda params, indices, axis=0, name=None: _w(np.take(np.asarray(params), np.asarray(indices), axis=axis))  # noqa: E731
squeeze = lambda x, axis=None, name=None: np.squeeze(x, axis=tuple(axis) if isinstance(axis, (list, tuple)) else axis)  # noqa: E731
expand_dims = lambda x, axis, name=None: np.expand_dims(x, axis)  # noqa: E731
square = _f(np.square)
sqrt = _f(np.sqrt)
exp = _f(np.exp)
sin = _f(np.sin)
tanh = _f(np.tanh)
sign = _f(np.sign)
abs = _f(np.abs)  # noqa: A001
rsqrt = lambda x, name=None: (1 / np.sqrt(np.asarray(x))).astype(np.asarray(x).dtype)  # noqa: E731
sigmoid = lambda x, name=None: (1 / (1 + np.exp(-np.asarray(x)))).astype(np.asarray(x).dtype)  # noqa: E731
softplus = lambda x, name=None: np.logaddexp(np.asarray(x), 0).astype(np.asarray(x).dtype)  # noqa: E731


_SEEDS = [1000]


class random_normal_initializer:  # noqa: N801
    """Draws from NumPy's generator, not TensorFlow's: same distribution, different numbers (the drivers overwrite
    trained variables with given arrays anyway)."""

    def __init__(self, mean=0.0, stddev=0.05, seed=None):
        self.mean, self.stddev, self.seed = mean, stddev, seed

    def __call__(self, shape, dtype=None, partition_info=None):
        seed = self.seed
        if seed is None:  # deterministic across runs of the generator
            _SEEDS[0] += 1
            seed = _SEEDS[0]
        return (np.random.default_rng(seed).standard_normal(tuple(shape)) * self.stddev + self.mean).astype(np.float32)


class constant_initializer:  # noqa: N801
    def __init__(self, value=0.0):
        self.value = value

    def __call__(self, shape, dtype=None, partition_info=None):
        return np.full(tuple(shape), self.value, np.float32)


def range(start, limit=None, delta=1, dtype=None, name=None):  # noqa: A001
    if limit is None:
        start, limit = 0, start
    out = np.arange(start, limit, delta)
    return out.astype(dtype.as_numpy_dtype) if dtype is not None else out.astype(np.int32)


def cumsum(x, axis=0, exclusive=False, reverse=False, name=None):
    x = np.asarray(x)
    if reverse:
        x = np.flip(x, axis)
    out = np.cumsum(x, axis=axis, dtype=x.dtype)
    if exclusive:
        out = out - x
    return _w(np.flip(out, axis) if reverse else out)


def mod(x, y, name=None):
    return np.mod(x, y)


def group(ops, name=None):
    return list(ops)


@contextlib.contextmanager
def control_dependencies(ops):
    yield


@contextlib.contextmanager
def name_scope(name):
    yield


math = types.SimpleNamespace(softplus=softplus, tanh=tanh, square=square, sigmoid=sigmoid, abs=abs, exp=exp, sin=sin,
                             sqrt=sqrt, rsqrt=rsqrt, log=_f(np.log), maximum=maximum, minimum=minimum)


# ---- compat.v1: variable scopes (names only), initializers, the global step -----------------------------
from . import _data  # noqa: E402

data = types.SimpleNamespace(Dataset=_data.Dataset, TFRecordDataset=_data.TFRecordDataset,
                             experimental=types.SimpleNamespace())
io = types.SimpleNamespace(gfile=types.SimpleNamespace(glob=_data.glob), FixedLenFeature=_data.FixedLenFeature,
                           VarLenFeature=_data.VarLenFeature, parse_single_example=_data.parse_single_example)
strings = types.SimpleNamespace(unicode_decode=_data.unicode_decode)
sparse = types.SimpleNamespace(to_dense=lambda x, **k: x)
FixedLenFeature, VarLenFeature, parse_single_example = _data.FixedLenFeature, _data.VarLenFeature, _data.parse_single_example
string = DType("string", np.object_, False)


def function(fn=None, **kwargs):
    """tf.function: the stand-in runs Python eagerly."""
    return fn if fn is not None else (lambda f: f)


def convert_to_tensor(x, dtype=None, name=None):
    return np.asarray(x) if dtype is None else cast(np.asarray(x), dtype)


class _Scope:
    def __init__(self):
        self.stack = []

    @property
    def name(self):
        return "/".join(self.stack)


_SCOPE = _Scope()


@contextlib.contextmanager
def _variable_scope(name, *a, **k):
    if isinstance(name, _Scope):  # re-entering a captured scope object: the stand-in has one scope stack
        yield _SCOPE
        return
    _SCOPE.stack.append(name)
    try:
        yield _SCOPE
    finally:
        _SCOPE.stack.pop()


class _ZerosInitializer:
    def __call__(self, shape, dtype=None, partition_info=None):
        return np.zeros(shape, np.float32)


GLOBAL_STEP = [np.asarray(0, np.int64)]
_train = types.SimpleNamespace(get_or_create_global_step=lambda: GLOBAL_STEP[0])


class _V1:
    """tf.compat.v1: its own few names, everything else falls through to the top-level stand-ins."""
    variable_scope = staticmethod(_variable_scope)
    get_variable_scope = staticmethod(lambda: _SCOPE)
    zeros_initializer = _ZerosInitializer
    train = _train
    logging = types.SimpleNamespace(info=lambda *a, **k: None)
    data = data

    def __getattr__(self, name):
        try:
            return globals()[name]
        except KeyError:
            raise AttributeError(f"tf.compat.v1.{name} is not part of the stand-in") from None


compat = types.SimpleNamespace(v1=_V1())


def __getattr__(name):
    if name[:1].isupper():
        cls = type(name, (), {})
        globals()[name] = cls
        return cls
    raise AttributeError(f"tensorflow.{name} is not part of the stand-in (oracle/ref_shim/README.md)")
