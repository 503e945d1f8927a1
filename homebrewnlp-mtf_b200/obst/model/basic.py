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
    keep = 1
    for extra in args.name_extras:
        if isinstance(extra, str) and extra.startswith('dropout_rate'):
            keep = 1 - float(extra[len('dropout_rate'):])
    if keep != 1:
        raise NotImplementedError("dropout with keep_prob < 1 is not on the hot path (all configs use rate 0)")
    return args.tensor


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


rezero = _unsupported('rezero')
group_linear = _unsupported('group_linear')
sum_heads = _unsupported('sum_heads')
transpose_sequence_features = _unsupported('transpose_sequence_features')
reduced_half_linear = _unsupported('reduced_half_linear')
product_key_memory = _unsupported('product_key_memory')
feed_forward_product_key_memory = _unsupported('feed_forward_product_key_memory')
