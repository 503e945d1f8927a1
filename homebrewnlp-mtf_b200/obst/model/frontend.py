"""Layer-string grammar and registry (src/model/frontend.py:21-75): a block is a list of strings
"name-extra1-extra2..."; each is dispatched through LAYER_FUNCTIONS with a BlockArgs carrying the extras."""
from ..dataclass import BlockArgs, BlockConfig, ModelParameter
from ..utils_core import scoped
from ..utils_mtf import Tensor
from .activation import activate
from .basic import (bottleneck_group_linear, dropout, feed_forward, feed_forward_product_key_memory, group_linear,
                    product_key_memory, reduced_half_linear, rezero, sum_heads, transpose_sequence_features)
from .normalization import norm
from .spatial import attention, cummean, cumsum


def _get_block_part(block_part_config: BlockConfig, params: ModelParameter, block_input: Tensor) -> Tensor:
    skip = block_part_config.skip and block_part_config.memory_reduction_strategy in ("none", "checkpoint")
    if skip:  # out = f(x) + x: f's backward goes on a tape of its own, the skip entry adds the identity branch
        return _with_skip(block_part_config, params, block_input)
    return _run_layers(block_part_config, params, block_input)


def _with_skip(block_part_config: BlockConfig, params: ModelParameter, block_input: Tensor) -> Tensor:
    import torch

    from .. import native
    from .backend import TapeEntry, run_tape
    outer = params.tape
    params.tape = [] if outer is not None else None
    try:
        out = _run_layers(block_part_config, params, block_input)
    finally:
        sub, params.tape = params.tape, outer
    if tuple(out.data.shape) != tuple(block_input.data.shape):
        raise ValueError(f"skip connection needs matching shapes, got {out.dims} and {block_input.dims}")
    res = torch.empty_like(out.data)
    native.axpby(out.data, block_input.data, 1.0, res)
    if outer is not None:
        def bwd(dy, mask, acc):
            return run_tape(sub, dy, acc=dy.clone())  # dx = dy + J^T dy

        outer.append(TapeEntry(bwd))
    return Tensor(res, out.dims)


def _run_layers(block_part_config: BlockConfig, params: ModelParameter, block_input: Tensor) -> Tensor:
    out = block_input
    layers = list(block_part_config.layer)
    skip_next = False
    for idx, layer in enumerate(layers, 1):
        if skip_next:
            skip_next = False
            continue
        name, *extras = layer.split('-')
        args = BlockArgs(params, out, extras, idx == len(layers))
        # a residual buffer handed down by the reversible driver may only be consumed by the LAST layer
        pending = params.residual
        if idx != len(layers):
            params.residual = None
        if name == 'norm' and idx < len(layers) and layers[idx].split('-') == ['activation', 'gelu']:
            # peephole: "norm-..." directly followed by "activation-gelu" runs as one fused kernel
            out = scoped(name + '_', norm, args, None, True)
            skip_next = True
        else:
            out = scoped(name + '_', LAYER_FUNCTIONS[name], args)
        if idx != len(layers):
            params.residual = pending
    return out


def block_part_fn(params: ModelParameter, block_part_config: BlockConfig, block_input: Tensor,
                  name_prefix: str = 'block') -> Tensor:
    return scoped(f"{name_prefix}_", _get_block_part, block_part_config, params, block_input)


def split_path(args: BlockArgs) -> Tensor:
    """src/model/frontend.py:39-55: "split_path-add;<layers>;<layers>..." (or multiply): every ';'-separated path is a
    ','-separated list of layer strings applied to the SAME input; the path outputs are summed (or multiplied).
    Each path records its own tape; one composite entry sends dy (times the other paths' outputs, for multiply)
    down every path and sums the input gradients."""
    import torch

    from .. import native
    from .backend import TapeEntry, run_tape
    base, *paths = '-'.join(str(e) for e in args.name_extras).split(';')
    base = base.split('-')
    if 'add' in base:
        multiply = False
    elif 'multiply' in base:
        multiply = True
    else:
        raise ValueError
    params = args.params
    outer = params.tape
    pending, params.residual = params.residual, None  # a residual can only ride on a single final GEMM
    outs, tapes = [], []
    for conf in paths:
        params.tape = [] if outer is not None else None
        try:
            outs.append(_run_layers(BlockConfig({'skip': False, 'layer': conf.split(',')}, ''), params, args.tensor))
        finally:
            tapes.append(params.tape)
            params.tape = outer
    params.residual = pending
    if not outs:
        raise ValueError("split_path needs at least one path")
    total = outs[0].data.clone()
    for o in outs[1:]:
        if multiply:
            native.mul(total, o.data, total)
        else:
            native.axpby(total, o.data, 1.0, total)
    if outer is not None:
        def bwd(dy, mask, acc):
            dx = acc
            for i, tape in enumerate(tapes):
                g = dy
                if multiply:  # d/d(path i) of the product = dy * product of the other paths
                    g = dy.clone()
                    for j, o in enumerate(outs):
                        if j != i:
                            native.mul(g, o.data, g)
                if dx is None:
                    dx = torch.zeros_like(args.tensor.data)
                dx = run_tape(tape, g, acc=dx)
            return dx

        outer.append(TapeEntry(bwd, supports_acc=True))
    return Tensor(total, outs[0].dims)


def _unsupported(name):
    def fn(args: BlockArgs) -> Tensor:
        raise NotImplementedError(f"layer '{name}' is registered by the reference but is outside the mixer hot path")

    return fn


LAYER_FUNCTIONS = {'feed_forward': feed_forward,
                   'attention': attention,
                   'cummean': cummean,
                   'cumsum': cumsum,
                   'norm': norm,
                   'rezero': rezero,
                   'activation': activate,
                   'convolution': _unsupported('convolution'),
                   'dropout': dropout,
                   'group_linear': group_linear,
                   'split_path': split_path,
                   'feed_forward_product_key_memory': feed_forward_product_key_memory,
                   'product_key_memory': product_key_memory,
                   'reduced_half_linear': reduced_half_linear,
                   'transpose_sequence_features': transpose_sequence_features,
                   'bottleneck_group_linear': bottleneck_group_linear,
                   'sum_heads': sum_heads
                   }
