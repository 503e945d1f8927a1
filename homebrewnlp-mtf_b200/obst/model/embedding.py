"""Embeddings (src/model/embedding.py): `embed` creates the absolute (learned) embedding variable -- used for the
token table, the LM head ("absolute-orthogonal") and the [heads, seq, seq] token-mixing maps -- or one of the
composed forms (sinusoidal `relative`, `relative-learned`, `axial`); `gather_embed` is the Gather / ScatterAdd op pair
(src/model/embedding.py:39-125)."""
import math
import typing

import torch

from .. import native
from ..dataclass import BlockArgs
from ..utils_core import scoped
from ..utils_mtf import DIM_LIST, Tensor
from .backend import TapeEntry, Variable, normal_var, orthogonal_var


def _embed_var(args: BlockArgs, shape: DIM_LIST) -> Variable:
    if 'orthogonal' in args:
        return orthogonal_var(args, shape)  # no fan-in dims: QR of a single column = one unit vector
    return normal_var(args, shape, args.params.embedding_stddev)


class Constant:
    """A tensor the graph reads like a variable but that is not trained (the sinusoidal relative embedding)."""
    trainable = False
    grad = None

    def __init__(self, value: torch.Tensor, dims: DIM_LIST):
        self.value = value
        self.data = value
        self.dims = list(dims)

    def ensure_grad(self):
        return None

    def set_masked(self, s: int):
        if s:
            raise NotImplementedError("a constant embedding cannot carry the causal mask")


def _multi_dim_range(dims: DIM_LIST, dtype: torch.dtype, device) -> torch.Tensor:
    """_multi_dim_range_tf (src/model/embedding.py:16-22): sum over dims of index * (product of the earlier sizes),
    cast to the activation dtype."""
    out = None
    stride = 1
    for idx, d in enumerate(dims):
        shape = [1] * idx + [d.size] + [1] * (len(dims) - idx - 1)
        part = (torch.arange(d.size, dtype=torch.float32, device=device) * stride).view(shape)
        out = part if out is None else out + part
        stride *= d.size
    return out.to(dtype)


def relative_embedding(args: BlockArgs, shape: DIM_LIST) -> torch.Tensor:
    """RelativeEmbeddingForward.lower (src/model/embedding.py:128-172): sin(position * exp(feature schedule)) *
    embedding_stddev, no gradient.  Built once per call site with torch (an initialisation-time constant, evaluated
    op by op in the activation dtype like the TF ops of the reference)."""
    params = args.params
    dt = params.variable_dtype.activation_dtype
    dev = params.device
    position_dims = [d for d in shape if d not in params.feature_dims and d not in params.intermediate]
    feature_dims = [d for d in shape if d not in position_dims]
    position_count = 1
    for d in position_dims:
        position_count *= d.size
    feature_count = 1
    for d in feature_dims:
        feature_count *= d.size
    positions = _multi_dim_range(position_dims, dt, dev)
    features = _multi_dim_range(feature_dims, dt, dev)
    additive = 0
    if 'cosine' in params.position_embedding:
        additive = torch.remainder(features, 2)
        features = (features - additive) / 2
        additive = additive * math.pi
        feature_count /= 2
    features = features + 4 / feature_count
    features = features - math.log(position_count / 2 / math.pi)
    features = torch.exp(features) + additive
    pos_shape = [d.size if d in position_dims else 1 for d in shape]
    feat_shape = [d.size if d in feature_dims else 1 for d in shape]
    out = positions.reshape(pos_shape) * features.reshape(feat_shape)
    return (torch.sin(out) * params.embedding_stddev).to(dt).contiguous()


class Composite:
    """An embedding that is a FUNCTION of trained variables (the axial factors' outer product, the learned scale of a
    relative embedding): consumers read `.value` and add into `.ensure_grad()` exactly as with a variable; a
    pass-through tape entry recorded when the embedding is built -- so it runs after every consumer's backward --
    carries that gradient on to the variables behind it."""
    trainable = False

    def __init__(self, args: BlockArgs, dims: DIM_LIST, value: torch.Tensor, backward):
        self.dims = list(dims)
        self.value = value.contiguous()
        self.data = self.value
        self.grad: typing.Optional[torch.Tensor] = None
        self._backward = backward
        if args.params.tape is not None:
            args.params.tape.append(TapeEntry(self._bwd))

    def ensure_grad(self) -> torch.Tensor:
        if self.grad is None:
            self.grad = torch.zeros(self.value.shape, dtype=torch.float32, device=self.value.device)
        return self.grad

    def _bwd(self, dy, mask, acc):
        if self.grad is not None:
            self._backward(self.grad)
            self.grad = None
        return dy  # not a layer: the activation gradient passes through untouched

    def set_masked(self, s: int):
        if s:
            raise NotImplementedError("a composed (axial / relative-learned) embedding cannot carry the causal mask")

    def transposed(self, *view):
        raise NotImplementedError("a composed (axial / relative-learned) embedding has no transposed shadow")


def _split_shape(args: BlockArgs, shape: DIM_LIST) -> typing.Tuple[DIM_LIST, DIM_LIST]:
    """position dims and feature dims of an embedding shape (src/model/embedding.py:186-188); the composed forms
    need shape == position dims + feature dims, which is how every caller builds it ([dim] + feature_dims)."""
    from ..utils_mtf import linear_shapes
    params = args.params
    position_dims = [d for d in shape if d not in params.feature_dims and d not in params.intermediate]
    feature_dims = list(linear_shapes(args).old)
    if list(shape) != position_dims + feature_dims:
        raise NotImplementedError(f"composed embedding of shape {shape}: expected {position_dims + feature_dims}")
    return position_dims, feature_dims


def _axial(args: BlockArgs, shape: DIM_LIST) -> Composite:
    """src/model/embedding.py:192-212: every position dim is factored into `splits` parts (the first takes what the
    integer root leaves over), each part owns a [part, features...] variable, and the embedding is their outer
    product over the parts, broadcast-multiplied over the features."""
    from ..utils_mtf import Dimension
    position_dims, feature_dims = _split_shape(args, shape)
    splits = 2
    for a in args:
        if isinstance(a, str) and a.isdigit():
            splits = int(a)
            break
    parts: typing.List[Variable] = []
    sizes: typing.List[int] = []
    for dim in position_dims:
        base = int(dim.size ** (1 / splits))
        while dim.size % base != 0:
            base -= 1
        for size in [dim.size // base ** (splits - 1)] + [base] * (splits - 1):
            parts.append(_embed_var(args, [Dimension(f"_{len(sizes)}", size)] + feature_dims))
            sizes.append(size)
    feat = [d.size for d in feature_dims]
    n = len(sizes)

    def view(i: int, t: torch.Tensor) -> torch.Tensor:  # part i broadcast against the others
        return t.float().view([sizes[j] if j == i else 1 for j in range(n)] + feat)

    out = None
    for i, var in enumerate(parts):
        out = view(i, var.value) if out is None else out * view(i, var.value)
    full = [d.size for d in shape]

    def backward(grad: torch.Tensor):
        g = grad.view(sizes + feat)
        for i, var in enumerate(parts):
            others = None
            for j, other in enumerate(parts):
                if j != i:
                    others = view(j, other.value) if others is None else others * view(j, other.value)
            gi = g if others is None else g * others
            reduce = [j for j in range(n) if j != i]
            var.ensure_grad().add_((gi.sum(dim=reduce) if reduce else gi).view(var.ensure_grad().shape))

    return Composite(args, shape, out.reshape(full).to(parts[0].value.dtype), backward)


def _relative_learned(args: BlockArgs, shape: DIM_LIST) -> Composite:
    """src/model/embedding.py:214-217: the sinusoidal embedding times a learned [features...] scale."""
    position_dims, feature_dims = _split_shape(args, shape)
    const = relative_embedding(args, shape)
    scale = _embed_var(args, feature_dims)
    value = (const.float() * scale.value.float()).to(const.dtype)

    def backward(grad: torch.Tensor):
        g = (grad * const.float()).sum(dim=list(range(len(position_dims))))
        scale.ensure_grad().add_(g.view(scale.ensure_grad().shape))

    return Composite(args, shape, value, backward)


def _embed(args: BlockArgs, shape: DIM_LIST):
    shape = list(shape)
    if 'absolute' in args:
        return _embed_var(args, shape)
    if 'axial' in args:
        return _axial(args, shape)
    if 'relative' in args:
        if 'learned' in args:
            return _relative_learned(args, shape)
        return Constant(relative_embedding(args, shape), shape)
    raise ValueError("The following embeddings are supported:"
                     " relative(-learned) or absolute(-split) or axial(-split) are supported")


def embed(args: BlockArgs, shape: DIM_LIST) -> Variable:
    return scoped('embed', _embed, args, shape)


def gather_embed(args: BlockArgs, shape: DIM_LIST, squeezed_dims: typing.Optional[DIM_LIST] = None) -> Tensor:
    """out[..., :] = table[idx[...], :]; backward = scatter-add into the table's fp32 gradient."""
    table = scoped("gather", embed, args, shape)
    idx = args.tensor
    width = shape[-1].size
    out = torch.empty(list(idx.data.shape) + [width], dtype=table.value.dtype, device=idx.data.device)
    native.embed_gather(idx.data, table.value, out)
    if args.params.tape is not None:
        def bwd(dy, mask, acc):
            native.embed_scatter_add(idx.data, dy, table.ensure_grad())
            return None

        args.params.tape.append(TapeEntry(bwd))
    args.params.tensor_storage.text_input_embedding = table
    return Tensor(out, idx.dims + [shape[-1]])
