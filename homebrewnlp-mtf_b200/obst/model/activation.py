"""Activations (src/model/activation.py:158-211): `activate` picks the first extra that names a known function and
falls back to identity with the reference's message.  relu and the reference's non-standard gelu have kernels;
the other registered names raise NotImplementedError (SURVEY.md 8(f) "next")."""
import torch

from .. import native
from ..dataclass import BlockArgs
from ..utils_core import scoped
from ..utils_mtf import Tensor
from .backend import TapeEntry


def _pointwise(kind: int):
    def fn(args: BlockArgs) -> Tensor:
        x = args.tensor.data
        y = torch.empty_like(x)
        native.act_fwd(kind, x, y)
        if args.params.tape is not None:
            def bwd(dy, mask, acc):
                dx = torch.empty_like(x)
                native.act_bwd(kind, x, dy, dx)
                return dx

            args.params.tape.append(TapeEntry(bwd, out_relu=y if kind == 0 else None))
        return Tensor(y, args.tensor.dims)

    return fn


def _unsupported(name: str):
    def fn(args: BlockArgs) -> Tensor:
        raise NotImplementedError(f"activation '{name}' is registered by the reference but has no B200 kernel yet")

    return fn


relu = _pointwise(0)
gelu = _pointwise(1)

ACTIVATIONS = {'relu': relu,
               'gelu': gelu,
               'sigmoid': _unsupported('sigmoid'),
               'tanh': _unsupported('tanh'),
               'lecun_tanh': _unsupported('lecun_tanh'),
               'silu': _unsupported('silu'),
               'mish': _unsupported('mish'),
               'mtf_mish': _unsupported('mtf_mish'),
               'softsign': _unsupported('softsign'),
               'exp': _unsupported('exp')}


_WARNED = set()


def activation_name(args: BlockArgs):
    for fn_name in args:
        if fn_name in ACTIVATIONS:
            return fn_name
    return None


def activate(args: BlockArgs) -> Tensor:
    """Call activation function on the tensor (first matching extra wins)."""
    for fn_name in args:
        if fn_name not in ACTIVATIONS:
            continue
        return scoped(fn_name, ACTIVATIONS[fn_name], args)
    key = tuple(str(e) for e in args.name_extras)
    if key not in _WARNED:  # the reference prints this while BUILDING the graph, i.e. once per call site
        _WARNED.add(key)
        print(f'No activation function found for "{args.name_extras}". Falling back to identity. '
              f'Known functions: {list(ACTIVATIONS.keys())}')
    return args.tensor
