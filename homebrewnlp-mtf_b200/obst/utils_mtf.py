"""Named-dimension algebra that decides every weight shape on the path (src/utils_mtf.py:84-120,189-195,
354-426 of the reference), restated on a minimal Dimension/Tensor pair instead of mesh_tensorflow types."""
import typing

import torch

from .utils_core import random_name  # noqa: F401  (re-exported like the reference module does)


class Dimension(typing.NamedTuple):
    name: str
    size: int


DIM_LIST = typing.List[Dimension]
LINEAR_SHAPES = typing.NamedTuple("LinearShapes", (("old", DIM_LIST), ("new", DIM_LIST)))
ATTENTION_DIM = typing.NamedTuple("AttentionDim", (("index", int), ("dim", Dimension)))


class Tensor:
    """A device buffer plus the names of its dimensions (what mtf.Tensor carries on this path)."""

    __slots__ = ("data", "dims")

    def __init__(self, data: torch.Tensor, dims: typing.Sequence[Dimension]):
        dims = list(dims)
        if tuple(d.size for d in dims) != tuple(data.shape):
            raise ValueError(f"shape mismatch: {tuple(data.shape)} vs {dims}")
        self.data = data
        self.dims = dims

    @property
    def shape(self) -> DIM_LIST:
        return self.dims

    @property
    def size(self) -> int:
        return self.data.numel()

    @property
    def dtype(self):
        return self.data.dtype

    def __repr__(self):
        return f"Tensor({[(d.name, d.size) for d in self.dims]}, {self.data.dtype})"


def dims_from_shape(shape) -> DIM_LIST:
    if isinstance(shape, Tensor):
        return list(shape.dims)
    return list(shape)


def deduplicate(inp):
    return type(inp)(dict.fromkeys(list(inp)))


def shape_sub(a, b) -> DIM_LIST:
    """mtf.Shape.__sub__: dims of a not in b."""
    b = dims_from_shape(b)
    return [d for d in dims_from_shape(a) if d not in b]


def shape_size(shape) -> int:
    out = 1
    for d in dims_from_shape(shape):
        out *= d.size
    return out


def dim_name(dim: typing.Union[Dimension, str]) -> str:
    return dim.name if isinstance(dim, Dimension) else dim


def anonymize_dim(dim: Dimension, new_size: typing.Optional[int] = None) -> Dimension:
    """leading underscore = 'anonymous' copy of a dim (src/utils_mtf.py:84-99)."""
    name = dim_name(dim)
    if not name.startswith("_"):
        name = "_" + name
    return Dimension(name, dim.size if new_size is None else new_size)


def unanonymize_dim(dim: Dimension, new_size: typing.Optional[int] = None) -> Dimension:
    name = dim_name(dim)
    if name.startswith("_"):
        name = name[1:]
    return Dimension(name, dim.size if new_size is None else new_size)


def shape_addition(*shapes) -> DIM_LIST:
    dims = []
    for s in shapes:
        dims.extend(dims_from_shape(s))
    return deduplicate(dims)


def shape_crossection(*shapes) -> DIM_LIST:
    shapes = [dims_from_shape(s) for s in shapes]
    return [dim for dim in shape_addition(*shapes) if all(dim in shape for shape in shapes)]


def get_intermediate(args) -> DIM_LIST:
    """src/utils_mtf.py:376-380"""
    if "group" not in args:
        return list(args.params.intermediate)
    return [args.params.head_dim,
            anonymize_dim(args.params.key_dim, args.params.key_dim.size * args.params.group_linear_factor)]


def linear_shapes(args) -> LINEAR_SHAPES:
    """src/utils_mtf.py:383-391: which dims a linear layer contracts ("old") and produces ("new")."""
    features = get_intermediate(args) + list(args.params.feature_dims)
    if "group" in args and args.params.intermediate[-1] in args.tensor.shape:
        features.remove(args.params.key_dim)
        features.extend(args.params.intermediate)
    features = deduplicate(features)
    old = shape_crossection(args.tensor.shape, features)
    drop = shape_sub(old, [args.params.head_dim] if "group" in args and args.params.head_dim in old else [])
    new = shape_sub(features, drop)
    return LINEAR_SHAPES(old, new)


def get_attention_dim(args) -> ATTENTION_DIM:
    """src/utils_mtf.py:418-422"""
    attention_dims = shape_sub(shape_sub(args.tensor.shape, args.params.feature_dims), args.params.intermediate)[1:]
    idx = args.params.attention_idx % len(attention_dims)
    return ATTENTION_DIM(idx, attention_dims[idx])


def is_masked(args) -> bool:
    """src/utils_mtf.py:425-426"""
    return get_attention_dim(args).index in args.params.masked_attention_dimensions


def feature_dims_used(params, shape, dims=None) -> bool:
    """src/utils_mtf.py:354-361"""
    shape = dims_from_shape(shape)
    if dims is None:
        dims = list(params.feature_dims) + [anonymize_dim(dim) for dim in params.feature_dims]
        return bool(sum(f in shape for f in dims) // 2)
    return all(f in shape for f in dims)
