"""NumPy stand-in for the slice of Mesh-TensorFlow the reference's Python touches (see ../README.md).

Semantics restated from mesh_tensorflow/ops.py (0.1.x): tensors carry NAMED dimensions; binary ops broadcast over the
union of the operands' dimensions (`_infer_binary_broadcast_shape`: the shape that contains the other, else shape1
followed by the dims only shape2 has); `einsum(xs, output_shape, reduced_dims)` contracts every dimension that is not
in the output; reductions name either the output shape or the reduced dimension; an Operation's name is prefixed with
the current variable scope.  A tensor's value is computed on first use from its operation's inputs, so
`Graph.clone_operations` (the reversible backward re-runs a block on reconstructed inputs) works as in the library;
every op carries the gradient rule of its mesh_tensorflow counterpart (EinsumOperation, AddOperation, Scalar*Operation,
the cwise grad_functions, ReduceOperation, reshape / broadcast / shift / slice), so the reference's own reverse sweep
(src/optimizer/__init__.py:127-174, src/model/revnet.py:55-120) runs on top of it."""
import builtins as _builtins
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
        return Shape(self._dims[key]) if isinstance(key, _builtins.slice) else self._dims[key]

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
utils = _Anything()  # utils.outside_all_rewrites is set at the end of this module

_DEFAULT_GRAPH = Graph()


def _cast_np(arr, dtype):
    return tf.cast(arr, dtype)


class _FakeMeshImpl:
    """lowering.mesh_impl(op): one 'processor', so slicewise(fn, *slices) is fn applied to the whole arrays."""

    def slicewise(self, fn, *inputs):
        return fn(*[np.asarray(x).view(tf.TFArray) for x in inputs])

    def import_tf_tensor(self, tensor, value):
        return value

    def laid_out_size(self, shape):
        return shape.size


class _FakeLowering:
    class _Values:
        def __getitem__(self, tensor):
            return tensor.np

    def __init__(self):
        self.tensors = self._Values()

    def mesh_impl(self, op):
        return _FakeMeshImpl()

    def set_tensor_lowering(self, tensor, value):
        tensor._value = _cast_np(np.asarray(value), tensor.dtype).reshape(tensor.shape.to_integer_list)


class Tensor:
    """Named-dimension tensor.  Its value (`.np`) is computed on first use by its operation: ops built by the
    stand-in evaluate a stored NumPy closure, Operation subclasses defined by the REFERENCE run their own lower()."""

    def __init__(self, operation, shape, dtype, name=None, index=0, value=None):
        self.operation = operation
        self.shape = convert_to_shape(shape)
        self.dtype = dtype
        self.name = name or (f"{operation.name}:{index}" if operation is not None else "tensor")
        self.usable = True
        self._value = None
        if value is not None:
            self.np = value

    @property
    def np(self):
        if self._value is None:
            self.operation._evaluate()
            if self._value is None:
                raise RuntimeError(f"{self.operation.name} did not produce {self.name}")
        return self._value

    @np.setter
    def np(self, value):
        self._value = _cast_np(np.asarray(value), self.dtype).reshape(self.shape.to_integer_list)

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
    def __floordiv__(self, o): return _binary(np.floor_divide, self, o, grad=None, name="floordiv")  # noqa: E704
    def __mod__(self, o): return _binary(np.mod, self, o, grad=None, name="mod")  # noqa: E704
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
        raise NotImplementedError(f"{type(self).__name__} has no gradient in the stand-in")

    def _evaluate(self):
        """An Operation subclass of the reference: run its own lower() on NumPy values."""
        self.lower(_FakeLowering())


class _Fn(Operation):
    """An op of the stand-in: `compute(op)` -> value from op.inputs[i].np (so a clone with other inputs recomputes),
    `grad(op, dy)` -> one gradient (or None) per input, as mesh_tensorflow's gradient() methods define them."""

    def __init__(self, inputs, shape, dtype, compute, grad, name, mesh=None):
        super().__init__(inputs, mesh=mesh, name=name)
        self._compute = compute
        self._grad = grad
        self._outputs = [Tensor(self, shape, dtype)]

    has_gradient = property(lambda self: self._grad is not None)

    def gradient(self, grad_ys):
        return self._grad(self, grad_ys[0])

    def _evaluate(self):
        self._outputs[0].np = self._compute(self)


def _op(inputs, shape, dtype, compute, grad, name, mesh=None):
    return _Fn(inputs, shape, dtype, compute, grad, name, mesh).outputs[0]


class Variable(Operation):
    def __init__(self, mesh, name, shape, dtype, initializer, trainable, **kwargs):
        super().__init__([], mesh, name=name)
        self._trainable = trainable
        self._shape = convert_to_shape(shape)
        self._dtype = dtype
        self.initializer = initializer
        ints = self._shape.to_integer_list
        init = initializer(ints, dtype.slice_dtype) if callable(initializer) else initializer
        if trainable and VARIABLE_VALUES:  # the driver imposes the weights, in creation order
            init = np.asarray(VARIABLE_VALUES.pop(0), dtype=np.float32)
            if init.size != self._shape.size:
                raise ValueError(f"imposed value of {init.shape} does not fit variable {self.name} {self._shape}")
        self._outputs = [Tensor(self, self._shape, dtype.activation_dtype)]
        self.set(np.asarray(init, dtype=np.float32))

    trainable = property(lambda self: self._trainable)
    shape = property(lambda self: self._shape)
    size = property(lambda self: self._shape.size)
    dtype = property(lambda self: self._dtype)
    value = property(lambda self: self._outputs[0])
    activation_dtype = property(lambda self: self._dtype.activation_dtype)
    has_gradient = property(lambda self: False)

    def set(self, array):
        """Write the slice (master) value; the graph-visible output is its activation-dtype cast."""
        self.master = _cast_np(np.asarray(array).reshape(self._shape.to_integer_list), self._dtype.slice_dtype)
        self._outputs[0].np = self.master

    def _evaluate(self):
        self._outputs[0].np = self.master


VARIABLE_VALUES = []  # see Variable.__init__


def get_variable(mesh, name, shape, dtype=None, initializer=None, trainable=True, **kwargs):
    if dtype is None:
        dtype = VariableDType(tf.float32)
    elif isinstance(dtype, tf.DType):
        dtype = VariableDType(dtype, dtype, dtype)
    return Variable(mesh, name, convert_to_shape(shape), dtype, initializer, trainable, **kwargs).outputs[0]


def _clone_operations(self, ops, input_mapping):
    """Graph.clone_operations: shallow copies of `ops` with their inputs remapped and fresh (unevaluated) outputs;
    Variables are shared, not cloned.  Returns (new ops, mapping old tensor -> new tensor)."""
    import copy as _copy
    mapping = dict(input_mapping)
    out = []
    for op in ops:
        if isinstance(op, Variable):
            continue
        new = _copy.copy(op)
        new._inputs = [mapping.get(t, t) for t in op._inputs]
        new._outputs = [Tensor(new, t.shape, t.dtype, t.name, i) for i, t in enumerate(op._outputs)]
        self.operations.append(new)
        for old, fresh in zip(op._outputs, new._outputs):
            mapping[old] = fresh
        out.append(new)
    return out, mapping


Graph.clone_operations = _clone_operations
Graph.combine_assignments = lambda self, ops: ops


class Assign(Operation):
    """assign / assign_add / assign_sub: recorded, applied by `apply()` -- after every value of the step has been
    computed from the OLD variable values, which is what evaluating a TF graph of reads followed by assigns does."""

    def __init__(self, variables, new_vals, fn, name):
        super().__init__(new_vals, variables[0].mesh, name=name)
        self.variables = variables
        self.new_vals = new_vals
        self.fn = fn
        self._staged = None

    has_gradient = property(lambda self: False)

    def stage(self):
        self._staged = [self.fn(var.master.astype(np.float32), _align_np(val.np, val.shape, var.shape).astype(np.float32))
                        for var, val in zip(self.variables, self.new_vals)]

    def apply(self):
        if self._staged is None:
            self.stage()
        for var, new in zip(self.variables, self._staged):
            var.set(new)
        self._staged = None

    def _evaluate(self):
        pass


def _vars(v):
    return [v] if isinstance(v, Variable) else list(v)


def assign(var, new_val, assign_fn=None, name=None):
    return Assign(_vars(var), [new_val] if isinstance(new_val, Tensor) else list(new_val), lambda old, new: new, "assign")


def assign_add(var, new_val, name=None):
    return Assign(_vars(var), [new_val], lambda old, new: old + new, "assign_add")


def assign_sub(var, new_val, name=None):
    return Assign(_vars(var), [new_val], lambda old, new: old - new, "assign_sub")


# ---- named-dimension arithmetic ----------------------------------------------------------------------------
def _align_np(arr, have_shape, shape):
    """arr (laid out as have_shape) laid out along `shape`; dims it lacks become size-1 axes."""
    shape = convert_to_shape(shape)
    have = have_shape.dimension_names
    for d in have_shape.dims:
        if d.name not in shape.dimension_names:
            raise ValueError(f"{have_shape} does not fit into {shape}")
    order = [have.index(d.name) for d in shape.dims if d.name in have]
    arr = np.transpose(arr, order) if order else arr
    return arr.reshape([d.size if d.name in have else 1 for d in shape.dims])


def _broadcast_shape(s1, s2):
    if all(d in s1.dims for d in s2.dims):
        return s1
    if all(d in s2.dims for d in s1.dims):
        return s2
    return Shape(s1.dims + [d for d in s2.dims if d not in s1.dims])


def _f32(arr, dtype):
    """Operand in the precision the op computes in: float32 for float32 / bfloat16 tensors (bf16 results are rounded
    afterwards), float64 for float64 tensors."""
    if not dtype.is_floating:
        return arr
    return arr.astype(np.float64 if dtype.name == "float64" else np.float32)


def _sum_to(dy, shape):
    """gradient of a broadcast: sum dy down to `shape`."""
    shape = convert_to_shape(shape)
    return dy if dy.shape == shape else reduce_sum(dy, output_shape=shape)


def _scalar(x):
    return not isinstance(x, Tensor)


def _binary(fn, x1, x2, output_shape=None, out_dtype=None, grad="none", name=None):
    """Binary op with broadcasting over named dims; a Python scalar operand stays a closure constant."""
    name = name or fn.__name__
    if _scalar(x1):
        x1, x2, fn = x2, x1, (lambda a, b, f=fn: f(b, a))
        swapped = True
    else:
        swapped = False
    if _scalar(x2):
        c = x2
        x = x1
        shape = convert_to_shape(output_shape) or x.shape
        g = None if grad in (None, "none") else (lambda op, dy: [grad(op, dy, c, swapped)])
        return _op([x], shape, out_dtype or x.dtype,
                   lambda op: np.broadcast_to(fn(_align_np(_f32(op.inputs[0].np, x.dtype), x.shape, shape),
                                                 np.asarray(c, dtype=x.dtype.as_numpy_dtype)), shape.to_integer_list),
                   g, "scalar_" + name)
    shape = convert_to_shape(output_shape) or _broadcast_shape(x1.shape, x2.shape)
    s1, s2 = x1.shape, x2.shape
    g = None if grad in (None, "none") else (lambda op, dy: grad(op, dy))
    return _op([x1, x2], shape, out_dtype or x1.dtype,
               lambda op: np.broadcast_to(fn(_align_np(_f32(op.inputs[0].np, x1.dtype), s1, shape),
                                             _align_np(_f32(op.inputs[1].np, x2.dtype), s2, shape)),
                                          shape.to_integer_list), g, name)


def add(x1, x2, output_shape=None, name=None):
    if _scalar(x2) or _scalar(x1):  # ScalarAddOperation: d/dx = dy
        return _binary(np.add, x1, x2, output_shape, grad=lambda op, dy, c, sw: dy, name="add")
    return _binary(np.add, x1, x2, output_shape,
                   grad=lambda op, dy: [_sum_to(dy, op.inputs[0].shape), _sum_to(dy, op.inputs[1].shape)], name="add")


def negative(x, name=None):
    return multiply(x, -1)


def sub(x1, x2, output_shape=None, name=None):
    if _scalar(x2):
        return add(x1, -x2, output_shape)
    if _scalar(x1):
        return add(negative(x2), x1, output_shape)
    return add(x1, negative(x2), output_shape)


def multiply(x1, x2, output_shape=None, name=None):
    if _scalar(x2) or _scalar(x1):  # ScalarMultiplyOperation: d/dx = dy * c
        return _binary(np.multiply, x1, x2, output_shape, grad=lambda op, dy, c, sw: dy * c, name="mul")
    return einsum([x1, x2], output_shape=convert_to_shape(output_shape) or _broadcast_shape(x1.shape, x2.shape))


def divide(x1, x2, output_shape=None, name=None):
    if _scalar(x2):
        return multiply(x1, 1.0 / x2)
    if _scalar(x1):
        return multiply(reciprocal(x2), x1)
    return multiply(x1, reciprocal(x2), output_shape)


def _minmax(fn, pick):
    def op(x1, x2, output_shape=None, name=None):
        def grad_tt(o, dy):
            a, b = o.inputs
            mask = cast(pick(a, b), dy.dtype)
            return [_sum_to(dy * mask, a.shape), _sum_to(dy * (1 - mask), b.shape)]

        def grad_ts(o, dy, c, swapped):
            return dy * cast(pick(o.inputs[0], c), dy.dtype)
        if _scalar(x1) or _scalar(x2):
            return _binary(fn, x1, x2, output_shape, grad=grad_ts, name=fn.__name__)
        return _binary(fn, x1, x2, output_shape, grad=grad_tt, name=fn.__name__)
    return op


def _compare(fn):
    def op(x1, x2, output_shape=None, name=None):
        return _binary(fn, x1, x2, output_shape, out_dtype=tf.bool, grad=None, name=fn.__name__)
    return op


equal, not_equal = _compare(np.equal), _compare(np.not_equal)
less, less_equal = _compare(np.less), _compare(np.less_equal)
greater, greater_equal = _compare(np.greater), _compare(np.greater_equal)
minimum = _minmax(np.minimum, lambda a, b: less_equal(a, b))
maximum = _minmax(np.maximum, lambda a, b: greater_equal(a, b))
logical_and = lambda a, b, output_shape=None, name=None: _binary(np.logical_and, a, b, output_shape, out_dtype=tf.bool, grad=None)  # noqa: E731


def logical_not(x, name=None):
    return _op([x], x.shape, tf.bool, lambda op: np.logical_not(op.inputs[0].np), None, "logical_not")


def add_n(xs):
    xs = list(xs)
    out = xs[0]
    for x in xs[1:]:
        out = add(out, x)
    return out


def _unary(fn, x, grad, name):
    return _op([x], x.shape, x.dtype, lambda op: fn(_f32(op.inputs[0].np, x.dtype)), grad, name)


# cwise ops and their mesh_tensorflow grad functions (ops.py: "grad_function=lambda op, dy: ...")
def square(x, name=None): return _unary(np.square, x, lambda op, dy: [dy * op.inputs[0] * 2], "square")  # noqa: E704
def sqrt(x, name=None): return _unary(np.sqrt, x, lambda op, dy: [dy * 0.5 / op.outputs[0]], "sqrt")  # noqa: E704
def exp(x, name=None): return _unary(np.exp, x, lambda op, dy: [dy * op.outputs[0]], "exp")  # noqa: E704
def log(x, name=None): return _unary(np.log, x, lambda op, dy: [dy / op.inputs[0]], "log")  # noqa: E704
def sin(x, name=None): return _unary(np.sin, x, lambda op, dy: [dy * cos(op.inputs[0])], "sin")  # noqa: E704
def cos(x, name=None): return _unary(np.cos, x, lambda op, dy: [dy * negative(sin(op.inputs[0]))], "cos")  # noqa: E704
def sign(x, name=None): return _unary(np.sign, x, None, "sign")  # noqa: E704
def abs(x, name=None): return _unary(np.abs, x, lambda op, dy: [dy * sign(op.inputs[0])], "abs")  # noqa: E704,A001


def rsqrt(x, name=None):
    return _unary(lambda v: 1 / np.sqrt(v), x,
                  lambda op, dy: [dy * -0.5 * (op.outputs[0] * op.outputs[0] * op.outputs[0])], "rsqrt")


def reciprocal(x, name=None):
    return _unary(lambda v: 1 / v, x, lambda op, dy: [negative(dy * square(op.outputs[0]))], "reciprocal")


def tanh(x, name=None):
    return _unary(np.tanh, x, lambda op, dy: [dy * (1.0 - square(op.outputs[0]))], "tanh")


def sigmoid(x, name=None):
    return _unary(lambda v: 1 / (1 + np.exp(-v)), x, lambda op, dy: [dy * op.outputs[0] * (1.0 - op.outputs[0])], "sigmoid")


def relu(x, name=None):
    return _unary(lambda v: np.maximum(v, 0), x, lambda op, dy: [dy * cast(greater(op.inputs[0], 0), dy.dtype)], "relu")


def softplus(x, name=None):
    return _unary(lambda v: np.logaddexp(v, 0), x, lambda op, dy: [dy * sigmoid(op.inputs[0])], "softplus")


def identity(x, name=None): return _unary(lambda v: v, x, lambda op, dy: [dy], "identity")  # noqa: E704
def stop_gradient(x, name=None): return _op([x], x.shape, x.dtype, lambda op: op.inputs[0].np, None, "stop_gradient")  # noqa: E704
def zeros_like(x): return _op([x], x.shape, x.dtype, lambda op: np.zeros(x.shape.to_integer_list), None, "zeros_like")  # noqa: E704
def ones_like(x): return _op([x], x.shape, x.dtype, lambda op: np.ones(x.shape.to_integer_list), None, "ones_like")  # noqa: E704


def pow(x, y):  # noqa: A001  -- mesh_tensorflow: exp(log(x) * y)
    return exp(log(x) * y)


def cast(x, dtype, name=None):
    if dtype == x.dtype:
        return x
    src = x.dtype
    grad = (lambda op, dy: [cast(dy, src)]) if (src.is_floating and dtype.is_floating) else None
    return _op([x], x.shape, dtype, lambda op: op.inputs[0].np, grad, "cast")


def constant(mesh, value, shape=None, dtype=tf.float32, name=None):
    shape = convert_to_shape(shape or [])
    return _op([], shape, dtype, lambda op: np.broadcast_to(np.asarray(value, dtype=dtype.as_numpy_dtype),
                                                            shape.to_integer_list), None, "constant", mesh=mesh)


def zeros(mesh, shape, dtype=tf.float32, name=None):
    return constant(mesh, 0, shape, dtype)


def ones(mesh, shape, dtype=tf.float32, name=None):
    return constant(mesh, 1, shape, dtype)


_NP_TO_TF = {np.dtype(np.float32): tf.float32, np.dtype(np.int32): tf.int32, np.dtype(np.int64): tf.int64,
             np.dtype(np.float64): tf.float64, np.dtype(np.bool_): tf.bool}


def import_fully_replicated(mesh, tf_tensor, shape, name=None):
    arr = np.asarray(tf_tensor)
    return _op([], shape, _NP_TO_TF[arr.dtype], lambda op: arr, None, name or "import", mesh=mesh)


import_tf_tensor = import_fully_replicated


def import_laid_out_tensor(mesh, laid_out_tensor, shape, name=None):
    return import_fully_replicated(mesh, laid_out_tensor, shape, name)


def broadcast(x, new_shape):
    new_shape = convert_to_shape(new_shape)
    src = x.shape
    return _op([x], new_shape, x.dtype,
               lambda op: np.broadcast_to(_align_np(op.inputs[0].np, src, new_shape), new_shape.to_integer_list),
               lambda op, dy: [_sum_to(dy, src)], "broadcast")


def reshape(x, new_shape, name=None):
    new_shape = convert_to_shape(new_shape)
    src = x.shape
    return _op([x], new_shape, x.dtype, lambda op: op.inputs[0].np.reshape(new_shape.to_integer_list),
               lambda op, dy: [reshape(dy, src)], "reshape")


def rename_dimension(x, old_name, new_name):
    return reshape(x, x.shape.rename_dimension(old_name, new_name))


def anonymize(x):
    return x


def transpose(x, new_shape, name=None):
    new_shape = convert_to_shape(new_shape)
    src = x.shape
    return _op([x], new_shape, x.dtype, lambda op: _align_np(op.inputs[0].np, src, new_shape),
               lambda op, dy: [transpose(dy, src)], "transpose")


def range(mesh, dim, dtype, name=None):  # noqa: A001
    dim = convert_to_dimension(dim)
    return _op([], [dim], dtype, lambda op: np.arange(dim.size).astype(dtype.as_numpy_dtype), None, "range", mesh=mesh)


def one_hot(indices, output_dim, on_value=1.0, off_value=0.0, dtype=tf.float32, name=None):
    return _op([indices], indices.shape + [output_dim], dtype,
               lambda op: np.where(np.equal(op.inputs[0].np[..., None], np.arange(output_dim.size)), on_value, off_value),
               None, "one_hot")


def _reduction(x, output_shape, reduced_dim):
    if reduced_dim is not None:
        if output_shape is not None:
            raise ValueError("set only one of output_shape and reduced_dim")
        if not isinstance(reduced_dim, Dimension) and isinstance(reduced_dim, (list, tuple, Shape)):
            reduced = [convert_to_dimension(d) for d in reduced_dim]
        else:
            reduced = [x.shape.get_dim_by_name(reduced_dim) if isinstance(reduced_dim, str) else convert_to_dimension(reduced_dim)]
        output_shape = x.shape - reduced
    output_shape = convert_to_shape(output_shape if output_shape is not None else [])
    axes = tuple(i for i, d in enumerate(x.shape.dims) if d not in output_shape.dims)
    kept = Shape([d for d in x.shape.dims if d in output_shape.dims])
    return output_shape, axes, kept


def reduce_sum(x, disable_positional_args=None, output_shape=None, reduced_dim=None, name=None):
    output_shape, axes, kept = _reduction(x, output_shape, reduced_dim)
    src = x.shape
    out = _op([x], kept, x.dtype, lambda op: np.sum(_f32(op.inputs[0].np, x.dtype), axis=axes),
              lambda op, dy: [broadcast(dy, src)], "reduce_sum")
    return out if kept == output_shape else transpose(out, output_shape)


def reduce_mean(x, disable_positional_args=None, output_shape=None, reduced_dim=None, name=None):
    out_shape, _, _ = _reduction(x, output_shape, reduced_dim)
    return reduce_sum(x, output_shape=out_shape) * (out_shape.size / x.shape.size)  # as mesh_tensorflow defines it


def _reduce_extreme(fn, name):
    def op(x, disable_positional_args=None, output_shape=None, reduced_dim=None, name_=None):
        output_shape, axes, kept = _reduction(x, output_shape, reduced_dim)

        def grad(o, dy):  # ReduceOperation.gradient for MAX / MIN: dy where the input attains the extreme
            return [dy * cast(equal(o.inputs[0], o.outputs[0]), dy.dtype)]
        out = _op([x], kept, x.dtype, lambda o: fn(o.inputs[0].np, axis=axes), grad, name)
        return out if kept == output_shape else transpose(out, output_shape)
    return op


reduce_max = _reduce_extreme(np.max, "reduce_max")
reduce_min = _reduce_extreme(np.min, "reduce_min")


def reduce_logsumexp(x, reduced_dim, extra_logit=None, name=None):
    reduced_dim = convert_to_dimension(reduced_dim)
    m = reduce_max(stop_gradient(x), reduced_dim=reduced_dim)
    return log(reduce_sum(exp(x - m), reduced_dim=reduced_dim)) + m


def einsum(xs, output_shape=None, reduced_dims=None, name=None):
    """EinsumOperation; gradient(i) = einsum([dy] + the other inputs, output_shape = input i's shape)."""
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
    letters = {d.name: chr(ord("a") + i)
               for i, d in enumerate(all_dims + [d for d in output_shape.dims if d not in all_dims])}
    shapes = [x.shape for x in xs]
    dts = [x.dtype for x in xs]
    spec = ",".join("".join(letters[d.name] for d in s.dims) for s in shapes)
    present = Shape([d for d in output_shape.dims if d in all_dims])
    spec += "->" + "".join(letters[d.name] for d in present.dims)

    def compute(op):
        value = np.einsum(spec, *[_f32(t.np, dt) for t, dt in zip(op.inputs, dts)])
        return np.broadcast_to(_align_np(value, present, output_shape), output_shape.to_integer_list)

    def grad(op, dy):
        ins = op.inputs
        return [einsum([dy] + [t for j, t in enumerate(ins) if j != i], output_shape=ins[i].shape) for i in builtins_range(len(ins))]

    return _op(xs, output_shape, xs[0].dtype, compute, grad, "einsum")


def shift(x, offset, dim, wrap, name=None):
    dim = x.shape.get_dim_by_name(dim) if isinstance(dim, str) else convert_to_dimension(dim)
    axis = x.shape.dims.index(dim)

    def compute(op):
        value = np.roll(op.inputs[0].np, offset, axis=axis)
        if not wrap and offset != 0:
            idx = [np.s_[:]] * value.ndim
            idx[axis] = np.s_[0:offset] if offset > 0 else np.s_[offset:]
            value = value.copy()
            value[tuple(idx)] = 0
        return value

    return _op([x], x.shape, x.dtype, compute, lambda op, dy: [shift(dy, -offset, dim, wrap)], "shift")


def pad(x, paddings, dim_name, name=None):
    axis = x.shape.dimension_names.index(dim_name)
    before, after = paddings
    new_shape = x.shape.resize_dimension(dim_name, x.shape.dims[axis].size + before + after)
    size = x.shape.dims[axis].size

    def compute(op):
        widths = [(0, 0)] * op.inputs[0].np.ndim
        widths[axis] = (before, after)
        return np.pad(op.inputs[0].np, widths)

    return _op([x], new_shape, x.dtype, compute, lambda op, dy: [slice(dy, before, size, dim_name)], "pad")


def slice(x, begin, size, slice_dim_name, name=None):  # noqa: A001
    axis = x.shape.dimension_names.index(slice_dim_name)
    total = x.shape.dims[axis].size

    def compute(op):
        idx = [np.s_[:]] * op.inputs[0].np.ndim
        idx[axis] = np.s_[begin:begin + size]
        return op.inputs[0].np[tuple(idx)]

    return _op([x], x.shape.resize_dimension(slice_dim_name, size), x.dtype, compute,
               lambda op, dy: [pad(dy, [begin, total - size - begin], slice_dim_name)], "slice")


def concat(xs, concat_dim_name, name=None):
    xs = list(xs)
    axis = xs[0].shape.dimension_names.index(concat_dim_name)
    sizes = [x.shape.dims[axis].size for x in xs]
    starts = [sum(sizes[:i]) for i in builtins_range(len(xs))]

    def grad(op, dy):
        return [slice(dy, b, s, concat_dim_name) for b, s in zip(starts, sizes)]

    return _op(xs, xs[0].shape.resize_dimension(concat_dim_name, sum(sizes)), xs[0].dtype,
               lambda op: np.concatenate([t.np for t in op.inputs], axis=axis), grad, "concat")


def argmax(x, reduced_dim, dtype=tf.int32, name=None):
    reduced_dim = convert_to_dimension(reduced_dim)
    axis = x.shape.dims.index(reduced_dim)
    return _op([x], x.shape - [reduced_dim], dtype, lambda op: np.argmax(op.inputs[0].np, axis=axis), None, "argmax")


def dropout(x, is_training, keep_prob=None, rate=None, noise_shape=None, name=None):
    rate = rate if rate is not None else (0.0 if keep_prob is None else 1 - keep_prob)
    if not is_training or rate == 0:
        return x
    raise NotImplementedError("random dropout masks are not reproduced by the stand-in")


def cwise(tf_fn, xs, output_dtype=None, grad_function=None, name=None):
    xs = list(xs)
    grad = None if grad_function is None else (lambda op, dy: grad_function(op, dy))
    return _op(xs, xs[0].shape, output_dtype or xs[0].dtype, lambda op: tf_fn(*[t.np for t in op.inputs]), grad,
               name or "cwise")


def slicewise(tf_fn, xs, output_shape=None, output_dtype=None, splittable_dims=None, grad_function=None, name=None):
    xs = list(xs)
    grad = None if grad_function is None else (lambda op, dy: grad_function(op, dy))
    return _op(xs, convert_to_shape(output_shape) or xs[0].shape, output_dtype or xs[0].dtype,
               lambda op: tf_fn(*[t.np for t in op.inputs]), grad, name or "slicewise")


UNIFORM_SOURCE = [None]  # the driver may supply (shape, minval, maxval) -> array, to share the draws with a checker


def random_uniform(mesh, shape, **kwargs):
    shape = convert_to_shape(shape)
    dtype = kwargs.get("dtype", tf.float32)
    lo, hi = kwargs.get("minval", 0), kwargs.get("maxval", 1)

    def compute(op):
        if UNIFORM_SOURCE[0] is not None:
            return UNIFORM_SOURCE[0](shape.to_integer_list, lo, hi)
        return np.random.default_rng(0).uniform(lo, hi, shape.to_integer_list)

    return _op([], shape, dtype, compute, None, "random_uniform", mesh=mesh)


BEFORE_LOOP_BODY = [None]  # the library traces a loop body ONCE; here it is re-run per iteration, and the driver uses
#                             this hook to rewind the reference's name counters so that every re-run finds the same names


def while_loop(cond_fn, body_fn, inputs, num_loop_vars=None, has_accumulators=False, **kwargs):
    """Evaluated on the spot: the loop runs while cond_fn's (scalar) value is true."""
    state = list(inputs)
    while bool(np.asarray(cond_fn(*state).np).reshape(-1)[0]):
        if BEFORE_LOOP_BODY[0] is not None:
            BEFORE_LOOP_BODY[0]()
        out = body_fn(*state)
        for t in out:
            if isinstance(t, Tensor):
                t.np  # noqa: B018 -- evaluate now: the next iteration must not re-run this one lazily
        state = list(out) + state[len(out):]
    return state


def recompute_grad(fn, explicit_inputs):
    return fn(*explicit_inputs)


builtins_range = _builtins.range


class _Rewrites:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


utils.outside_all_rewrites = lambda: _Rewrites()


def __getattr__(name):
    """Class names that only appear in annotations / isinstance checks (mtf.AddOperation, mtf.LazyAllreduceSum ...)."""
    if name[:1].isupper():
        cls = type(name, (Operation,), {})
        globals()[name] = cls
        return cls
    raise AttributeError(f"mesh_tensorflow.{name} is not part of the stand-in (oracle/ref_shim/README.md)")
