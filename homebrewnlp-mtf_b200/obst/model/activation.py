"""Activations (src/model/activation.py:158-211): `activate` picks the first extra that names a known function and
falls back to identity with the reference's message.  relu and the reference's non-standard gelu are the hot
path's (act_kernel, or fused into a GEMM epilogue / the norm kernel); the other registered names run through the
generic pointwise kernel, with the reference's own (non-analytic) backward for silu and mish."""
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


relu = _pointwise(0)
gelu = _pointwise(1)

# name -> OBST_ACT_* kind of include/obst_b200.h (src/model/activation.py:190-200)
ACTIVATION_KINDS = {'relu': 0, 'gelu': 1, 'sigmoid': 2, 'tanh': 3, 'lecun_tanh': 4, 'silu': 5, 'mish': 6,
                    'mtf_mish': 7, 'softsign': 8, 'exp': 9}
ACTIVATIONS = {name: _pointwise(kind) for name, kind in ACTIVATION_KINDS.items()}
ACTIVATIONS['relu'], ACTIVATIONS['gelu'] = relu, gelu


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
