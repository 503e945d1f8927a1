"""Eager NumPy stand-in for the slice of Mesh-TensorFlow the reference's Python touches (see ../README.md).

Semantics restated from mesh_tensorflow/ops.py (0.1.x): tensors carry NAMED dimensions; binary ops broadcast over the
union of the operands' dimensions (`_infer_binary_broadcast_shape`: the shape that contains the other, else shape1
followed by the dims only shape2 has); `einsum(xs, output_shape, reduced_dims)` contracts every dimension that is not
in the output; reductions name either the output shape or the reduced dimension; an Operation's name is prefixed with
the current variable scope.  Every op computes its value immediately (no graph lowering, no autodiff)."""
import collections

import numpy as np
import tensorflow as tf

Dimension = collections.namedtuple("Dimension", ["name", "size"])


def convert_to_dimension(d):
    if d is None or isinstance(d, Dimension):
        return d
    name, size = d
    return Dimension(name, size)


class Shape:
    def __init__(self, dims):
        if isinstance(dims, Shape):
            dims = dims.dims
        self._dims = [convert_to_dimension(d) for d in dims]
        names = [d.name for d in self._dims]
        if len(set(names)) != len(names):
            raise ValueError(f"Shape must not have repeated dimensions: {names}")

    dims = property(lambda self: list(self._dims))
    ndims = property(lambda self: len(self._dims))
    dimension_names = property(lambda self: [d.name for d in self._dims])
    to_integer_list = property(lambda self: [d.size for d in self._dims])

    @property
    def size(self):
        out = 1
        for d in self._dims:
            out *= d.size
        return out

    def __iter__(self):
        return iter(self._dims)

    def __len__(self):
        return len(self._dims)

    def __getitem__(self, key):
        return Shape(self._dims[key]) if isinstance(key, slice) else self._dims[key]

    def __eq__(self, other):
        return isinstance(other, (Shape, list, tuple)) and list(self._dims) == list(convert_to_shape(other)._dims)

    def __ne__(self, other):
        return not self == other

    def __hash__(self):
        return hash(tuple(self._dims))

    def __add__(self, other):
        if isinstance(other, Dimension):
            other = [other]
        return Shape(self._dims + list(convert_to_shape(other)._dims))

    def __sub__(self, other):
        if other is None:
            return self
        if isinstance(other, Dimension):
            other = [other]
        other = convert_to_shape(other)
        return Shape([d for d in self._dims if d not in other._dims])

    def __contains__(self, d):
        return d in self._dims

    def __repr__(self):
        return "Shape[%s]" % ", ".join(f"{d.name}={d.size}" for d in self._dims)

    to_string = property(__repr__)

    def get_dim_by_name(self, name):
        for d in self._dims:
            if d.name == name:
                return d
        raise ValueError(f"Dimension {name} not found in {self}")

    def rename_dimension(self, old, new):
        return Shape([Dimension(new, d.size) if d.name == old else d for d in self._dims])

    def resize_dimension(self, name, size):
        return Shape([Dimension(name, size) if d.name == name else d for d in self._dims])


def convert_to_shape(x):
    if x is None:
        return None
    return x if isinstance(x, Shape) else Shape(list(x))


class VariableDType:
    def __init__(self, master_dtype=tf.float32, slice_dtype=None, activation_dtype=None):
        self.master_dtype = master_dtype
        self.slice_dtype = slice_dtype or master_dtype
        self.activation_dtype = activation_dtype or master_dtype


class Graph:
    def __init__(self):
        self.operations = []
        self.name_to_variable = {}
        self.captured_variable_scope = tf.compat.v1.get_variable_scope()

    @property
    def trainable_variables(self):
        return [op for op in self.operations if isinstance(op, Variable) and op.trainable]

    @property
    def all_variables(self):
        return [op for op in self.operations if isinstance(op, Variable)]


class Mesh:
    def __init__(self, graph, name, variable_placer=None):
        self.graph = graph
        self.name = name


class Lowering:
    pass


class _Anything:
    """Namespace whose attributes are all classes: annotations such as mtf.simd_mesh_impl.SimdMeshImpl resolve."""

    def __getattr__(self, name):
        return type(name, (), {})


simd_mesh_impl = _Anything()
placement_mesh_impl = _Anything()
utils = _Anything()

_DEFAULT_GRAPH = Graph()


def _cast_np(arr, dtype):
    return tf.cast(arr, dtype)


class Tensor:
    def __init__(self, operation, shape, dtype, name=None, index=0, value=None):
        self.operation = operation
        self.shape = convert_to_shape(shape)
        self.dtype = dtype
        self.name = name or (f"{operation.name}:{index}" if operation is not None else "tensor")
        self.usable = True
        self.np = None if value is None else _cast_np(np.asarray(value), dtype).reshape(self.shape.to_integer_list)

    graph = property(lambda self: self.operation.graph)
    mesh = property(lambda self: self.operation.mesh)
    size = property(lambda self: self.shape.size)

    def dimension(self, name):
        return self.shape.get_dim_by_name(name)

    def __repr__(self):
        return f"Tensor[{self.name}, {self.shape}, {self.dtype}]"

    def __add__(self, o): return add(self, o)  # noqa: E704
    def __radd__(self, o): return add(self, o)  # noqa: E704
    def __sub__(self, o): return sub(self, o)  # noqa: E704
    def __rsub__(self, o): return add(negative(self), o)  # noqa: E704
    def __mul__(self, o): return multiply(self, o)  # noqa: E704
    def __rmul__(self, o): return multiply(self, o)  # noqa: E704
    def __neg__(self): return negative(self)  # noqa: E704
    def __truediv__(self, o): return divide(self, o)  # noqa: E704
    def __rtruediv__(self, o): return multiply(reciprocal(self), o)  # noqa: E704
    def __floordiv__(self, o): return _binary(np.floor_divide, self, o)  # noqa: E704
    def __mod__(self, o): return _binary(np.mod, self, o)  # noqa: E704
    def __pow__(self, o): return pow(self, o)  # noqa: E704,A001


class Operation:
    def __init__(self, inputs, mesh=None, name=None):
        if mesh is None:
            if not inputs:
                raise ValueError("mesh must be specified if no inputs")
            mesh = inputs[0].mesh
        self._inputs = list(inputs)
        self._outputs = []
        self._mesh = mesh
        scope = tf.compat.v1.get_variable_scope().name
        name = name or type(self).__name__
        self._name = f"{scope}/{name}" if scope else name
        mesh.graph.operations.append(self)

    inputs = property(lambda self: list(self._inputs))
    outputs = property(lambda self: list(self._outputs))
    mesh = property(lambda self: self._mesh)
    graph = property(lambda self: self._mesh.graph)
    name = property(lambda self: self._name)
    has_gradient = property(lambda self: True)

    def gradient(self, grad_ys):
        raise NotImplementedError("the stand-in evaluates forward arithmetic only (oracle/ref_shim/README.md)")


class _Eager(Operation):
    """Generic op whose single output was computed on the spot."""

    def __init__(self, inputs, value, shape, dtype, name, mesh=None):
        super().__init__([t for t in inputs if isinstance(t, Tensor)], mesh=mesh, name=name)
        self._outputs = [Tensor(self, shape, dtype, value=value)]


def _out(inputs, value, shape, dtype, name, mesh=None):
    return _Eager(inputs, value, shape, dtype, name, mesh).outputs[0]


class Variable(Operation):
    def __init__(self, mesh, name, shape, dtype, initializer, trainable, **kwargs):
        super().__init__([], mesh, name=name)
        self._trainable = trainable
        self._shape = convert_to_shape(shape)
        self._dtype = dtype
        self.initializer = initializer
        ints = self._shape.to_integer_list
        init = initializer(ints, dtype.slice_dtype) if callable(initializer) else initializer
        self.master = _cast_np(np.asarray(init, dtype=np.float32).reshape(ints), dtype.slice_dtype)
        self._outputs = [Tensor(self, self._shape, dtype.activation_dtype, value=self.master)]

    trainable = property(lambda self: self._trainable)
    shape = property(lambda self: self._shape)
    size = property(lambda self: self._shape.size)
    dtype = property(lambda self: self._dtype)
    value = property(lambda self: self._outputs[0])
    activation_dtype = property(lambda self: self._dtype.activation_dtype)

    def set(self, array):
        """Write the slice (master) value; the graph-visible output is its activation-dtype cast."""
        self.master = _cast_np(np.asarray(array).reshape(self._shape.to_integer_list), self._dtype.slice_dtype)
        self._outputs[0].np = _cast_np(self.master, self._dtype.activation_dtype)


def get_variable(mesh, name, shape, dtype=None, initializer=None, trainable=True, **kwargs):
    if dtype is None:
        dtype = VariableDType(tf.float32)
    elif isinstance(dtype, tf.DType):
        dtype = VariableDType(dtype, dtype, dtype)
    return Variable(mesh, name, convert_to_shape(shape), dtype, initializer, trainable, **kwargs).outputs[0]


class Assign(Operation):
    """assign / assign_add / assign_sub: recorded, applied by `apply()` -- after every value of the step has been
    computed from the OLD variable values, which is what evaluating a TF graph of reads followed by assigns does."""

    def __init__(self, variables, new_vals, fn, name):
        super().__init__(new_vals, variables[0].mesh, name=name)
        self.variables = variables
        self.new_vals = new_vals
        self.fn = fn

    def apply(self):
        for var, val in zip(self.variables, self.new_vals):
            new = _align(val, var.shape).astype(np.float32)
            var.set(self.fn(var.master.astype(np.float32), new))


def _vars(v):
    return [v] if isinstance(v, Variable) else list(v)


def assign(var, new_val, assign_fn=None, name=None):
    return Assign(_vars(var), [new_val] if isinstance(new_val, Tensor) else list(new_val), lambda old, new: new, "assign")


def assign_add(var, new_val, name=None):
    return Assign(_vars(var), [new_val], lambda old, new: old + new, "assign_add")


def assign_sub(var, new_val, name=None):
    return Assign(_vars(var), [new_val], lambda old, new: old - new, "assign_sub")


# ---- named-dimension arithmetic ----------------------------------------------------------------------------
def _align(t, shape):
    """t's value laid out along `shape` (dims t lacks become size-1 axes)."""
    shape = convert_to_shape(shape)
    have = t.shape.dimension_names
    for d in t.shape.dims:
        if d.name not in shape.dimension_names:
            raise ValueError(f"{t.shape} does not fit into {shape}")
    order = [have.index(d.name) for d in shape.dims if d.name in have]
    arr = np.transpose(t.np, order) if order else t.np
    return arr.reshape([d.size if d.name in have else 1 for d in shape.dims])


def _broadcast_shape(s1, s2):
    if all(d in s1.dims for d in s2.dims):
        return s1
    if all(d in s2.dims for d in s1.dims):
        return s2
    return Shape(s1.dims + [d for d in s2.dims if d not in s1.dims])


def _finish(value, dtype):
    return _cast_np(value, dtype)


def _to_tensor(x, like):
    if isinstance(x, Tensor):
        return x
    return _out([], np.asarray(x, dtype=like.dtype.as_numpy_dtype), [], like.dtype, "constant", mesh=like.mesh)


def _binary(fn, x1, x2, output_shape=None, out_dtype=None, name=None):
    if not isinstance(x1, Tensor):
        x1 = _to_tensor(x1, x2)
    x2 = _to_tensor(x2, x1)
    shape = convert_to_shape(output_shape) or _broadcast_shape(x1.shape, x2.shape)
    a, b = _align(x1, shape), _align(x2, shape)
    value = np.broadcast_to(fn(a, b), shape.to_integer_list)
    return _out([x1, x2], value, shape, out_dtype or x1.dtype, name or fn.__name__)


def _unary(fn, x, name=None):
    return _out([x], fn(x.np), x.shape, x.dtype, name or getattr(fn, "__name__", "unary"))


def add(x1, x2, output_shape=None, name=None): return _binary(np.add, x1, x2, output_shape)  # noqa: E704
def sub(x1, x2, output_shape=None, name=None): return _binary(np.subtract, x1, x2, output_shape)  # noqa: E704
def multiply(x1, x2, output_shape=None, name=None): return _binary(np.multiply, x1, x2, output_shape)  # noqa: E704
def divide(x1, x2, output_shape=None, name=None): return _binary(np.true_divide, x1, x2, output_shape)  # noqa: E704
def minimum(x1, x2, output_shape=None, name=None): return _binary(np.minimum, x1, x2, output_shape)  # noqa: E704
def maximum(x1, x2, output_shape=None, name=None): return _binary(np.maximum, x1, x2, output_shape)  # noqa: E704


def _compare(fn):
    def op(x1, x2, output_shape=None, name=None):
        return _binary(fn, x1, x2, output_shape, out_dtype=tf.bool)
    return op


equal, not_equal = _compare(np.equal), _compare(np.not_equal)
less, less_equal = _compare(np.less), _compare(np.less_equal)
greater, greater_equal = _compare(np.greater), _compare(np.greater_equal)
logical_and = lambda a, b, output_shape=None, name=None: _binary(np.logical_and, a, b, output_shape, out_dtype=tf.bool)  # noqa: E731
logical_not = lambda x, name=None: _out([x], np.logical_not(x.np), x.shape, tf.bool, "logical_not")  # noqa: E731


def add_n(xs):
    out = xs[0]
    for x in xs[1:]:
        out = add(out, x)
    return out


negative = lambda x, name=None: _unary(np.negative, x)  # noqa: E731
square = lambda x, name=None: _unary(np.square, x)  # noqa: E731
sqrt = lambda x, name=None: _unary(np.sqrt, x)  # noqa: E731
rsqrt = lambda x, name=None: _unary(lambda v: 1 / np.sqrt(v), x, "rsqrt")  # noqa: E731
reciprocal = lambda x, name=None: _unary(lambda v: 1 / v, x, "reciprocal")  # noqa: E731
exp = lambda x, name=None: _unary(np.exp, x)  # noqa: E731
log = lambda x, name=None: _unary(np.log, x)  # noqa: E731
sin = lambda x, name=None: _unary(np.sin, x)  # noqa: E731
tanh = lambda x, name=None: _unary(np.tanh, x)  # noqa: E731
sign = lambda x, name=None: _unary(np.sign, x)  # noqa: E731
relu = lambda x, name=None: _unary(lambda v: np.maximum(v, 0), x, "relu")  # noqa: E731
sigmoid = lambda x, name=None: _unary(lambda v: 1 / (1 + np.exp(-v)), x, "sigmoid")  # noqa: E731
softplus = lambda x, name=None: _unary(lambda v: np.logaddexp(v, 0), x, "softplus")  # noqa: E731
identity = lambda x, name=None: _unary(lambda v: v, x, "identity")  # noqa: E731
stop_gradient = lambda x, name=None: _unary(lambda v: v, x, "stop_gradient")  # noqa: E731
zeros_like = lambda x: _unary(np.zeros_like, x)  # noqa: E731
ones_like = lambda x: _unary(np.ones_like, x)  # noqa: E731


def pow(x, y):  # noqa: A001  -- mesh_tensorflow: exp(log(x) * y)
    return exp(log(x) * y)


def cast(x, dtype, name=None):
    if dtype == x.dtype:
        return x
    return _out([x], x.np, x.shape, dtype, "cast")


def constant(mesh, value, shape=None, dtype=tf.float32, name=None):
    shape = convert_to_shape(shape or [])
    return _out([], np.broadcast_to(np.asarray(value, dtype=dtype.as_numpy_dtype), shape.to_integer_list), shape, dtype,
                "constant", mesh=mesh)


def zeros(mesh, shape, dtype=tf.float32, name=None):
    return constant(mesh, 0, shape, dtype)


def ones(mesh, shape, dtype=tf.float32, name=None):
    return constant(mesh, 1, shape, dtype)


def import_fully_replicated(mesh, tf_tensor, shape, name=None):
    arr = np.asarray(tf_tensor)
    dtype = {np.dtype(np.float32): tf.float32, np.dtype(np.int32): tf.int32, np.dtype(np.int64): tf.int64,
             np.dtype(np.float64): tf.float64, np.dtype(np.bool_): tf.bool}[arr.dtype]
    return _out([], arr, shape, dtype, name or "import", mesh=mesh)


import_tf_tensor = import_fully_replicated


def broadcast(x, new_shape):
    new_shape = convert_to_shape(new_shape)
    return _out([x], np.broadcast_to(_align(x, new_shape), new_shape.to_integer_list), new_shape, x.dtype, "broadcast")


def reshape(x, new_shape, name=None):
    new_shape = convert_to_shape(new_shape)
    return _out([x], x.np.reshape(new_shape.to_integer_list), new_shape, x.dtype, "reshape")


def rename_dimension(x, old_name, new_name):
    return reshape(x, x.shape.rename_dimension(old_name, new_name))


def anonymize(x):
    return x


def transpose(x, new_shape, name=None):
    new_shape = convert_to_shape(new_shape)
    return _out([x], _align(x, new_shape), new_shape, x.dtype, "transpose")


def range(mesh, dim, dtype, name=None):  # noqa: A001
    dim = convert_to_dimension(dim)
    return _out([], np.arange(dim.size).astype(dtype.as_numpy_dtype), [dim], dtype, "range", mesh=mesh)


def one_hot(indices, output_dim, on_value=1.0, off_value=0.0, dtype=tf.float32, name=None):
    eye = np.equal(indices.np[..., None], np.arange(output_dim.size))
    value = np.where(eye, on_value, off_value)
    return _out([indices], value, indices.shape + [output_dim], dtype, "one_hot")


def _reduction_dims(x, output_shape, reduced_dim):
    if reduced_dim is not None:
        if output_shape is not None:
            raise ValueError("set only one of output_shape and reduced_dim")
        reduced_dim = convert_to_dimension(reduced_dim) if not isinstance(reduced_dim, str) else x.shape.get_dim_by_name(reduced_dim)
        output_shape = x.shape - [reduced_dim]
    output_shape = convert_to_shape(output_shape if output_shape is not None else [])
    axes = tuple(i for i, d in enumerate(x.shape.dims) if d not in output_shape.dims)
    kept = Shape([d for d in x.shape.dims if d in output_shape.dims])
    return output_shape, axes, kept


def _reduce(fn, name):
    def op(x, disable_positional_args=None, output_shape=None, reduced_dim=None, name_=None):
        output_shape, axes, kept = _reduction_dims(x, output_shape, reduced_dim)
        value = fn(x.np.astype(np.float32) if x.dtype.is_floating else x.np, axis=axes)
        tmp = _out([x], value, kept, x.dtype, name)
        return tmp if kept == output_shape else transpose(tmp, output_shape)
    return op


reduce_sum = _reduce(np.sum, "reduce_sum")
reduce_mean = _reduce(np.mean, "reduce_mean")
reduce_max = _reduce(np.max, "reduce_max")
reduce_min = _reduce(np.min, "reduce_min")


def reduce_logsumexp(x, reduced_dim, extra_logit=None, name=None):
    reduced_dim = convert_to_dimension(reduced_dim)
    m = reduce_max(stop_gradient(x), reduced_dim=reduced_dim)
    return log(reduce_sum(exp(x - m), reduced_dim=reduced_dim)) + m


def einsum(xs, output_shape=None, reduced_dims=None, name=None):
    xs = list(xs)
    all_dims = []
    for x in xs:
        for d in x.shape.dims:
            if d not in all_dims:
                all_dims.append(d)
    if output_shape is None:
        reduced = [convert_to_dimension(d) for d in (reduced_dims or [])]
        output_shape = Shape([d for d in all_dims if d not in reduced])
    output_shape = convert_to_shape(output_shape)
    letters = {d.name: chr(ord("a") + i) for i, d in enumerate(all_dims + [d for d in output_shape.dims if d not in all_dims])}
    spec = ",".join("".join(letters[d.name] for d in x.shape.dims) for x in xs)
    present = [d for d in output_shape.dims if d in all_dims]
    value = np.einsum(spec + "->" + "".join(letters[d.name] for d in present), *[x.np.astype(np.float32) if x.dtype.is_floating else x.np for x in xs])
    tmp = _out(xs, value, present, xs[0].dtype, "einsum")
    return tmp if Shape(present) == output_shape else broadcast(tmp, output_shape)


def shift(x, offset, dim, wrap, name=None):
    dim = convert_to_dimension(dim) if not isinstance(dim, str) else x.shape.get_dim_by_name(dim)
    axis = x.shape.dims.index(dim)
    value = np.roll(x.np, offset, axis=axis)
    if not wrap:
        idx = [slice(None)] * value.ndim
        idx[axis] = slice(0, offset) if offset > 0 else slice(offset, None)
        value = value.copy()
        if offset != 0:
            value[tuple(idx)] = 0
    return _out([x], value, x.shape, x.dtype, "shift")


def slice(x, begin, size, slice_dim_name, name=None):  # noqa: A001
    axis = x.shape.dimension_names.index(slice_dim_name)
    idx = [np.s_[:]] * x.np.ndim
    idx[axis] = np.s_[begin:begin + size]
    return _out([x], x.np[tuple(idx)], x.shape.resize_dimension(slice_dim_name, size), x.dtype, "slice")


def concat(xs, concat_dim_name, name=None):
    axis = xs[0].shape.dimension_names.index(concat_dim_name)
    value = np.concatenate([x.np for x in xs], axis=axis)
    return _out(xs, value, xs[0].shape.resize_dimension(concat_dim_name, value.shape[axis]), xs[0].dtype, "concat")


def cumsum(x, dim, exclusive=False):
    dim = convert_to_dimension(dim)
    axis = x.shape.dims.index(dim)
    value = np.cumsum(x.np.astype(np.float32), axis=axis)
    if exclusive:
        value = value - x.np
    return _out([x], value, x.shape, x.dtype, "cumsum")


def argmax(x, reduced_dim, dtype=tf.int32, name=None):
    reduced_dim = convert_to_dimension(reduced_dim)
    axis = x.shape.dims.index(reduced_dim)
    return _out([x], np.argmax(x.np, axis=axis), x.shape - [reduced_dim], dtype, "argmax")


def dropout(x, is_training, keep_prob=None, rate=None, noise_shape=None, name=None):
    rate = rate if rate is not None else (0.0 if keep_prob is None else 1 - keep_prob)
    if not is_training or rate == 0:
        return x
    raise NotImplementedError("random dropout masks are not reproduced by the stand-in")


def cwise(tf_fn, xs, output_dtype=None, grad_function=None, name=None):
    value = tf_fn(*[x.np for x in xs])
    return _out(xs, value, xs[0].shape, output_dtype or xs[0].dtype, name or "cwise")


def slicewise(tf_fn, xs, output_shape=None, output_dtype=None, splittable_dims=None, grad_function=None, name=None):
    value = tf_fn(*[x.np for x in xs])
    return _out(xs, value, convert_to_shape(output_shape) or xs[0].shape, output_dtype or xs[0].dtype, name or "slicewise")


def recompute_grad(fn, explicit_inputs):
    return fn(*explicit_inputs)


def __getattr__(name):
    """Class names that only appear in annotations / isinstance checks (mtf.AddOperation, mtf.LazyAllreduceSum ...)."""
    if name[:1].isupper():
        cls = type(name, (Operation,), {})
        globals()[name] = cls
        return cls
    raise AttributeError(f"mesh_tensorflow.{name} is not part of the stand-in (oracle/ref_shim/README.md)")
