"""Feed-forward family (src/model/basic.py).  On the hot path: `bottleneck_group_linear` =
activated_linear('in:') -> activated_linear('mid:') -> activated_linear('out:') with the `group` extras added in
between (src/model/basic.py:122-126).  ReLU is folded into the GEMM epilogue; dropout with keep_prob 1 is the
identity (src/model/basic.py:25-30)."""
from ..dataclass import BlockArgs
from ..utils_mtf import Tensor, linear_shapes
from .activation import activate, activation_name
from .backend import linear
from .normalization import norm


def _unsupported(name: str):
    def fn(args: BlockArgs) -> Tensor:
        raise NotImplementedError(f"layer '{name}' is registered by the reference but is outside the mixer hot path "
                                  f"(SURVEY.md 8(f))")

    return fn


def dropout(args: BlockArgs) -> Tensor:
    """src/model/basic.py:25-30 -> mtf.dropout: identity unless training with keep_prob < 1, else inverted dropout.
    The mask comes from a counter-based generator keyed by (init_seed, step, call site), so the backward pass (and a
    reversible recompute of the same call site within the step) sees the same mask."""
    import torch

    from .. import native
    from ..utils_core import current_scope
    from .backend import TapeEntry
    keep = 1
    for extra in args.name_extras:
        if isinstance(extra, str) and extra.startswith('dropout_rate'):
            keep = 1 - float(extra[len('dropout_rate'):])
    params = args.params
    if keep == 1 or not params.train:
        return args.tensor
    import zlib
    seed = (zlib.crc32(current_scope().encode()) << 32) ^ (int(params.global_step) * 2654435761 + int(params.init_seed))
    x = args.tensor.data
    out = torch.empty_like(x)
    native.dropout(x, out, keep, seed)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            dx = torch.empty_like(dy)
            native.dropout(dy, dx, keep, seed)
            return dx

        params.tape.append(TapeEntry(bwd))
    return Tensor(out, args.tensor.dims)


def wrapped_linear(args: BlockArgs, relu: bool = False) -> Tensor:
    return linear(args, *linear_shapes(args), relu=relu)


def activated_linear(args: BlockArgs, prefix: str) -> Tensor:
    """src/model/basic.py:47-57: linear -> activation -> dropout, optionally gated by sigmoid(second linear)
    (`glu`), plus activation(third linear) (`glu_add`), optionally normalised."""
    args = args([a[len(prefix):] for a in args if isinstance(a, str) and a.startswith(prefix)])
    if 'mixture_of_experts' in args:
        raise NotImplementedError("mixture_of_experts is outside the mixer hot path")
    act = activation_name(args)
    params = args.params
    gated = 'glu' in args or 'glu_add' in args
    pending = params.residual  # only a plain final linear may fold the residual into its epilogue
    if pending is not None and (act is not None or 'norm' in args or gated):
        params.residual = None

    def main_branch():
        if act == 'relu':  # fused: linear -> relu in one GEMM epilogue
            return dropout(args(wrapped_linear(args, relu=True)))
        return dropout(args(activate(args(wrapped_linear(args)))))

    if not gated:
        out = main_branch()
    else:
        out = _glu(args, main_branch, 'glu_add' in args)
    if 'norm' in args:
        out = norm(args(out))
    if pending is not None and params.residual is None and (act is not None or 'norm' in args or gated):
        params.residual = pending  # not consumed: the reversible driver adds it explicitly
    return out


def _glu(args: BlockArgs, main_branch, add: bool) -> Tensor:
    """out = main * sigmoid(linear_2(x)) [+ activation(linear_3(x))]: three branches off the same input, each
    recorded on its own tape; one composite entry sums the three input gradients."""
    import torch
    from .. import native
    from .backend import TapeEntry, run_tape
    params = args.params
    outer = params.tape

    def branch(fn):
        params.tape = [] if outer is not None else None
        y = fn()
        tape, params.tape = params.tape, outer
        return y, tape

    a, tape_a = branch(main_branch)
    b, tape_b = branch(lambda: wrapped_linear(args))
    out = torch.empty_like(a.data)
    native.gate_fwd(a.data, b.data, out)
    c, tape_c = (None, None)
    if add:
        c, tape_c = branch(lambda: activate(args(wrapped_linear(args))))
        native.axpby(out, c.data, 1.0, out)
    if outer is not None:
        def into(tape, g):
            # a branch that ends in a ReLU fused into its GEMM leaves the [y > 0] mask to whoever produces dy
            if tape and tape[-1].out_relu is not None:
                masked = torch.empty_like(g)
                native.act_bwd(0, tape[-1].out_relu, g, masked)
                return masked
            return g

        def bwd(dy, mask, acc):
            da, db = torch.empty_like(a.data), torch.empty_like(b.data)
            native.gate_bwd(a.data, b.data, dy, da, db)
            dx = run_tape(tape_a, into(tape_a, da), acc=acc)
            dx = run_tape(tape_b, db, acc=dx)
            if add:
                dx = run_tape(tape_c, into(tape_c, dy), acc=dx)
            return dx

        outer.append(TapeEntry(bwd, supports_acc=True))
    return Tensor(out, a.dims)


def activated_linear_in(args: BlockArgs) -> Tensor:
    return activated_linear(args, 'in:')


def activated_linear_out(args: BlockArgs) -> Tensor:
    return activated_linear(args, 'out:')


def feed_forward(args: BlockArgs) -> Tensor:
    pending, args.params.residual = args.params.residual, None
    args = args(activated_linear_in(args))
    args.params.residual = pending
    return activated_linear_out(args)


def bottleneck_group_linear(args: BlockArgs) -> Tensor:
    pending, args.params.residual = args.params.residual, None
    args = args(activated_linear_in(args))
    args.name_extras.extend(['group', 'mid:group', 'out:group'])
    args = args(activated_linear(args, 'mid:'))
    args.params.residual = pending  # y1 = x1 + f(x2): the add rides on the last (grouped) GEMM
    return activated_linear_out(args)


def rezero(args: BlockArgs) -> Tensor:
    """src/model/basic.py:21-22: x * s with s a rank-0 variable initialised to 0 (the optimizer gives rank-0 variables
    adam instead of sm3 and rezero variables their own learning-rate multiplier, src/optimizer/optimizers.py:61-62,
    src/optimizer/__init__.py:46-47)."""
    import numpy as np
    import torch

    from .. import native
    from .backend import TapeEntry, get_var
    var = get_var(args, [], lambda seed: np.zeros((), np.float32))
    x = args.tensor.data
    out = torch.empty_like(x)
    native.scale(x, var.value, out)
    if args.params.tape is not None:
        def bwd(dy, mask, acc):
            native.dot_accumulate(x, dy, var.ensure_grad())
            dx = torch.empty_like(dy)
            native.scale(dy, var.value, dx)
            return dx

        args.params.tape.append(TapeEntry(bwd))
    return Tensor(out, args.tensor.dims)


def group_linear(args: BlockArgs) -> Tensor:
    """src/model/basic.py:72-74: a per-head linear over the features of each head, reshaped back to the input's
    dims.  (`args('group')` in the reference appends the TYPE `str`, not the string -- BlockArgs.__call__ quirk -- so
    no `group` extra is actually added; nothing downstream of this call reads it.)"""
    from ..utils_mtf import anonymize_dim
    params = args.params
    anonymous_key = [anonymize_dim(d) if d == params.key_dim else d for d in params.feature_dims]
    out = linear(args('group'), params.feature_dims, anonymous_key)
    dims = [params.key_dim if d == anonymize_dim(params.key_dim) else d for d in out.dims]
    return Tensor(out.data, dims)


def sum_heads(args: BlockArgs) -> Tensor:
    """src/model/basic.py:77-78: reduce_sum over the head dim."""
    import torch

    from .. import native
    from .backend import TapeEntry
    params = args.params
    x = args.tensor
    if x.dims[-2:] != params.feature_dims:
        raise NotImplementedError(f"sum_heads: expected [..., heads, features_per_head], got {x.dims}")
    H, K = params.head_dim.size, params.key_dim.size
    rows = x.data.numel() // (H * K)
    out = torch.empty(list(x.data.shape[:-2]) + [K], dtype=x.data.dtype, device=x.data.device)
    native.sum_heads(x.data, out, rows, H, K)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            dx = torch.empty_like(x.data)
            native.broadcast_heads(dy, dx, rows, H, K)
            return dx

        params.tape.append(TapeEntry(bwd))
    return Tensor(out, x.dims[:-2] + [params.key_dim])


def transpose_sequence_features(args: BlockArgs) -> Tensor:
    """src/model/basic.py:81-86: exchange the sequence and features_per_head axes (they must have equal sizes)."""
    import torch

    from .. import native
    from .backend import TapeEntry
    params = args.params
    x = args.tensor
    assert params.features_per_head == params.sequence_length, "ToDo: Support other shapes"
    if len(x.dims) != 4 or x.dims[1] != params.sequence_dim or x.dims[2:] != params.feature_dims:
        raise NotImplementedError(f"transpose_sequence_features: expected [batch, sequence, heads, features], got {x.dims}")
    B, S, H, K = (d.size for d in x.dims)
    out = torch.empty_like(x.data)
    native.swap_axes(x.data, out, B, S, H, K)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            dx = torch.empty_like(dy)
            native.swap_axes(dy, dx, B, K, H, S)
            return dx

        params.tape.append(TapeEntry(bwd))
    return Tensor(out, x.dims)


def reduced_half_linear(args: BlockArgs) -> Tensor:
    """src/model/basic.py:89-90: group_linear of the head sum: every head's weight sees the same summed input."""
    return group_linear(args(sum_heads(args)))


product_key_memory = _unsupported('product_key_memory')
feed_forward_product_key_memory = _unsupported('feed_forward_product_key_memory')
