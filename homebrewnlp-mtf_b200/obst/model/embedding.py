"""Embeddings (src/model/embedding.py): `embed` creates the absolute (learned) embedding variable -- used for the
token table, the LM head ("absolute-orthogonal") and the [heads, seq, seq] token-mixing maps -- and
`gather_embed` is the Gather / ScatterAdd op pair (src/model/embedding.py:39-125)."""
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


def _embed(args: BlockArgs, shape: DIM_LIST):
    shape = list(shape)
    if 'absolute' in args:
        return _embed_var(args, shape)
    if 'axial' in args:
        raise NotImplementedError("axial embeddings are not implemented (SURVEY.md 8(f) item 3)")
    if 'relative' in args:
        if 'learned' in args:
            raise NotImplementedError("relative-learned embeddings are not implemented (SURVEY.md 8(f) item 3)")
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
