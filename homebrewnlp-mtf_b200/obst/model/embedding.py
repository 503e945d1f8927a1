"""Embeddings (src/model/embedding.py): `embed` creates the absolute (learned) embedding variable -- used for the
token table, the LM head ("absolute-orthogonal") and the [heads, seq, seq] token-mixing maps -- and
`gather_embed` is the Gather / ScatterAdd op pair (src/model/embedding.py:39-125)."""
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


def _embed(args: BlockArgs, shape: DIM_LIST) -> Variable:
    if 'absolute' in args:
        return _embed_var(args, list(shape))
    if 'axial' in args or 'relative' in args:
        raise NotImplementedError("axial / relative embeddings are outside the mixer hot path (SURVEY.md 8(f))")
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
