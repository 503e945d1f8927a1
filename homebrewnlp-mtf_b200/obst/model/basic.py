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
    if _can_fold_mid_norm(args):
        args.params.residual = pending
        return _mid_norm_out_folded(args)
    args = args(activated_linear(args, 'mid:'))
    args.params.residual = pending  # y1 = x1 + f(x2): the add rides on the last (grouped) GEMM
    return activated_linear_out(args)


def _sub_args(args: BlockArgs, prefix: str) -> BlockArgs:
    return args([a[len(prefix):] for a in args if isinstance(a, str) and a.startswith(prefix)])


def _can_fold_mid_norm(args: BlockArgs) -> bool:
    """mid = norm(relu(linear)) followed by out = plain grouped linear, bf16 on the tcgen05 path: the norm can be
    folded into the two GEMMs around it (see _mid_norm_out_folded)."""
    import torch
    params = args.params
    if not getattr(params, "fold_mid_norm", True) or params.gemm_backend == 1:
        return False
    if args.tensor.data.dtype != torch.bfloat16:
        return False
    mid, out = _sub_args(args, 'mid:'), _sub_args(args, 'out:')
    if activation_name(mid) != 'relu' or 'norm' not in mid or 'glu' in mid or 'glu_add' in mid \
            or 'mixture_of_experts' in mid:
        return False
    if activation_name(out) is not None or 'norm' in out or 'glu' in out or 'glu_add' in out \
            or 'mixture_of_experts' in out:
        return False
    for a in list(mid) + list(out):
        if isinstance(a, str) and a.startswith('dropout_rate') and float(a[len('dropout_rate'):]) != 0:
            return False
    k2 = params.key_dim.size * params.group_linear_factor
    return k2 % 256 == 0 and params.key_dim.size in (256, 512) and 'group' in mid


def _mid_norm_out_folded(args: BlockArgs) -> Tensor:
    """mid: h = relu(h1 @ W_mid), n = norm(h); out: f = n @ W_out per head (src/model/basic.py:47-57,122-126) --
    without ever writing n.  With n = (h - mu) r gamma + beta per (token, head):

        f       = r (h @ (gamma . W)) - r mu (gamma^T W) + beta^T W
        dW[j,k] = gamma_j ( sum_t h[t,j] r_t dy[t,k] - sum_t r_t mu_t dy[t,k] ) + beta_j sum_t dy[t,k]
        dn      = dy @ W^T, and the ordinary norm backward gives dh (it needs h and (mu, r) only)

    The statistics come out of W_mid's epilogue (obst_gemm epilogue 7), W_out's epilogue applies the norm to its
    accumulator together with the reversible add (5 / 6), the weight gradient's rows are scaled by gamma in its
    epilogue (8).  Per token and depth this saves the norm kernel's read of h and write of n, and W_out's read of n
    is replaced by the read of h: 32 KiB less HBM traffic per pass at 32big_mixer's shape.

    Inside the reversible backward the norm's own backward is folded as well: with
    s1 = dy . (gamma^T W) and s2 = dy . (h @ (gamma.W)) -- row dots with the accumulator of the recomputed W_out GEMM,
    epilogues 9 / 10 -- the dgrad epilogue (11) writes dh = [h > 0] (r gamma dn - r m1 - r xhat m2) directly, and
    dgamma / dbeta come out of the weight gradient's accumulator and two column sums; dn is never written."""
    import torch

    from .. import native
    from ..utils_mtf import shape_size, shape_sub
    from .backend import TapeEntry, orthogonal_var
    from .normalization import norm_layout
    params = args.params
    mid, out_args = _sub_args(args, 'mid:'), _sub_args(args, 'out:')
    pending, params.residual = params.residual, None
    h1 = mid.tensor
    T = shape_size(h1.dims[:-1])
    # ---- W_mid with the ReLU + statistics epilogue (variable order as in the unfused path: W_mid, gamma, beta, W_out)
    old, new = linear_shapes(mid)
    n_mid = shape_size(new)
    part = torch.empty((T, n_mid // 128, 2), dtype=torch.float32, device=h1.data.device)
    h = linear(mid, old, new, relu=True, relu_stats=part)
    heads, width, gamma, beta = norm_layout(mid(h))
    stats = torch.empty((T * heads, 2), dtype=torch.float32, device=h1.data.device)
    native.norm_stats_finalize(part, stats, T, heads, width // 128, width)
    del part
    # ---- W_out with the norm folded in
    oargs = out_args(h)
    old, new = linear_shapes(oargs)
    var = orthogonal_var(oargs, list(old) + list(new), list(old))
    group = [d for d in old if d in new]
    contract = [d for d in old if d not in group]
    produce = [d for d in new if d not in group]
    lead = shape_sub(h.shape, old)
    if group != [params.head_dim] or h.dims != lead + group + contract or shape_size(contract) != width:
        raise NotImplementedError(f"folded norm: unexpected layout {h.dims} / {old} -> {new}")
    G, Kc, Np = heads, width, shape_size(produce)
    wg_t, cvec, dvec, gamma_f, wg = var.fold_norm(gamma, beta, G, Kc, Np)
    out_dims = lead + group + produce
    if pending is not None:
        buf, sign, dst = pending
        if tuple(buf.shape) != tuple(d.size for d in out_dims):
            raise NotImplementedError("folded norm: residual of a different shape")
    else:  # no reversible driver around (layer-level use): add to zeros
        buf = dst = torch.zeros([d.size for d in out_dims], dtype=h.data.dtype, device=h.data.device)
        sign = 1.0
    hd = h.data
    # The reversible backward knows the gradient dy of this block's output while it recomputes the block
    # (params.upstream_grad, src/model/revnet.py:91-120): W_out's epilogue then also takes the two per-row reductions
    # of the norm's backward -- sum_j gamma_j dn_j = dy . (gamma^T W) and sum_j gamma_j h_j dn_j = dy . (h @ (gamma.W)),
    # the latter being a dot product with this very accumulator -- and dn = dy @ W^T never has to be written.
    up = getattr(params, "upstream_grad", None)
    # (the row-dot epilogues exist in the CTA-pair kernel only: more than one 128-row tile)
    fold_bwd = (T > 128 and getattr(params, "fold_mid_norm_backward", True) and up is not None and pending is not None and params.tape is not None and gamma is not None
                and up.dtype == hd.dtype and tuple(up.shape) == tuple(dst.shape) and up.is_contiguous())
    dots = None
    if fold_bwd:
        dots = torch.empty((T, G, 2 * ((Np + 255) // 256), 2), dtype=torch.float32, device=hd.device)
        native.gemm(native.operand(hd, G * Kc, 1, s_glo=Kc), native.operand(wg_t, Kc, 1, s_glo=Kc * Np), dst, T, Np, Kc,
                    GLO=G, d_ld=G * Np, d_glo=Np, R=buf, epi=9 if sign > 0 else 10, backend=2,
                    aux=(stats, cvec, dvec, up, dots))
    else:
        native.gemm(native.operand(hd, G * Kc, 1, s_glo=Kc), native.operand(wg_t, Kc, 1, s_glo=Kc * Np), dst, T, Np, Kc,
                    GLO=G, d_ld=G * Np, d_glo=Np, R=buf, epi=5 if sign > 0 else 6, backend=2, aux=(stats, cvec, dvec))
    if params.tape is not None:
        w = var.value
        g_val = None if gamma is None else gamma.value
        b_val = None if beta is None else beta.value

        def bwd(dy: torch.Tensor, mask, acc):
            # weight gradient from h and r . dy, rows scaled by gamma, plus the two rank-one terms
            dyr = torch.empty_like(dy)
            v = torch.zeros((2, G, Np), dtype=torch.float32, device=dy.device)
            native.rowscale_colsum(dy, stats, dyr, v[0], v[1], T, G, Np)
            grad = var.ensure_grad()
            folded = dots is not None and mask is not None and dy.data_ptr() == up.data_ptr()
            if folded:
                # ... the same GEMM also leaves sum_k W_jk G_jk in gamma's gradient (G = its raw accumulator), which
                # with the two column sums is the norm's scale / shift gradient
                dgamma = gamma.ensure_grad()
                dbeta = None if beta is None else beta.ensure_grad()
                native.gemm(native.operand(hd, 1, G * Kc, s_glo=Kc), native.operand(dyr, 1, G * Np, s_glo=Np), grad, Kc,
                            Np, T, GLO=G, d_ld=Np, d_glo=Kc * Np, accumulate=True, epi=8, backend=2,
                            aux=(gamma_f, w, dgamma))
                native.norm_affine_grad(w, v[0], v[1], dgamma, dbeta, G, Kc, Np)
                native.norm_wgrad_rank1(grad, g_val, b_val, v[0], v[1], G, Kc, Np)
                coeffs = torch.empty((T, G, 4), dtype=torch.float32, device=dy.device)
                native.norm_bwd_coeffs(dots, stats, coeffs, T, G, dots.shape[2], Kc)
                # dh = [h > 0] (a (dy @ (gamma.W)^T) + b h + c): the norm's and the ReLU's backward in the dgrad epilogue
                dh = torch.empty_like(hd)
                native.gemm(native.operand(dy, G * Np, 1, s_glo=Np), native.operand(wg, Np, 1, s_glo=Kc * Np), dh, T, Kc,
                            Np, GLO=G, d_ld=G * Kc, d_glo=Kc, R=hd, epi=11, backend=2, aux=(coeffs,))
                return dh
            native.gemm(native.operand(hd, 1, G * Kc, s_glo=Kc), native.operand(dyr, 1, G * Np, s_glo=Np), grad, Kc, Np,
                        T, GLO=G, d_ld=Np, d_glo=Kc * Np, accumulate=True, epi=8, backend=2, aux=(gamma_f,))
            native.norm_wgrad_rank1(grad, g_val, b_val, v[0], v[1], G, Kc, Np)
            # dn = dy @ W^T, then the norm's own backward (with the ReLU mask of h: it IS the norm's input)
            dn = torch.empty_like(hd)
            native.gemm(native.operand(dy, G * Np, 1, s_glo=Np), native.operand(w, Np, 1, s_glo=Kc * Np), dn, T, Kc, Np,
                        GLO=G, d_ld=G * Kc, d_glo=Kc, backend=params.gemm_backend)
            dh = torch.empty_like(hd)
            native.gn_bwd(dn, hd, g_val, b_val, dh, None if gamma is None else gamma.ensure_grad(),
                          None if beta is None else beta.ensure_grad(), G, Kc, relu_mask=mask is not None, stats=stats)
            return dh

        # `mask` is h itself (the previous entry's ReLU output): folded into the norm backward above
        params.tape.append(TapeEntry(bwd, supports_mask=True))
    return Tensor(dst, out_dims)


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
