"""Sequence-axis mixing (src/model/spatial.py:42-81).  On the hot path `attention` is called with
`biased_attention_map-absolute-input_as_value-shared`: no q/k/v, no softmax -- a learned causal [heads, seq, seq]
map applied to the input itself (MLP-Mixer token mixing):

    out[b,s,h,k] = sum_{t<=s} M[h,s,t] * x[b,t,h,k]

run as a batched lower-triangular tcgen05 GEMM that reads x in its [batch, seq, heads, features] layout."""
import torch

from .. import native
from ..dataclass import BlockArgs
from ..utils_mtf import Tensor, anonymize_dim, get_attention_dim, is_masked
from .backend import TapeEntry
from .embedding import embed


def attention(args: BlockArgs) -> Tensor:
    """src/model/spatial.py:42-81.  The mixer configs reach it as `biased_attention_map-absolute-input_as_value-
    shared` (a learned causal map applied to the input itself): that form runs as one causal GEMM (_map_mix).  Every
    other combination -- dot-product logits, `biased_softmax`, `scale_attention_map`, projected values -- goes through
    _general_attention."""
    params = args.params
    params.attention_idx += 1
    simple = ("dot_product" not in args and "input_as_value" in args and 'biased_softmax' not in args
              and 'scale_attention_map' not in args)
    if simple:
        if 'biased_attention_map' not in args:
            raise UserWarning(f"WARNING: There is no spatial mixing happening with the following attention parameters: "
                              f"{args.name_extras}.")
        return _map_mix(args)
    return _general_attention(args)


def _map_mix(args: BlockArgs) -> Tensor:
    params = args.params
    x = args.tensor
    dim = get_attention_dim(args).dim
    tmp = anonymize_dim(dim)
    if len(x.dims) != 4 or x.dims[1] != dim or x.dims[2:] != params.feature_dims:
        raise NotImplementedError(f"attention: expected [batch, {dim.name}, heads, features_per_head], got {x.dims}")
    var = embed(args, [params.head_dim, dim, tmp])  # `shared`: two maps in the whole model
    masked = is_masked(args)
    var.set_masked(dim.size if masked else 0)  # compare_range(..., greater_equal) folded into the weight shadow
    Bn, S, H, K = (d.size for d in x.dims)
    F = H * K
    be = params.gemm_backend
    dt = x.data.dtype
    xin = x.data
    w = var.value
    wt = var.transposed(H, S, S)  # M^T, so the dx GEMM reads its A operand K-major too
    residual = params.residual
    b_op = native.operand(xin, 1, F, s_glo=K, s_ghi=S * F)
    a_op = native.operand(w, S, 1, s_glo=S * S)
    if residual is not None:
        params.residual = None
        buf, sign, out = residual
        native.gemm(a_op, b_op, out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, R=buf,
                    epi=2 if sign > 0 else 3, tri=1 if masked else 0, dtype_in=dt, backend=be)
    else:
        out = torch.empty_like(xin)
        native.gemm(a_op, b_op, out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1 if masked else 0,
                    dtype_in=dt, backend=be)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            # dM[h,s,t] += [t<=s] sum_{b,k} dy[b,s,h,k] x[b,t,h,k]; the map is shared by every depth, so this
            # accumulates (fp32) across all of them
            native.gemm(native.operand(dy, F, 1, s_ko=S * F, s_glo=K), native.operand(xin, F, 1, s_ko=S * F, s_glo=K),
                        var.ensure_grad(), S, S, K, KO=Bn, GLO=H, d_ld=S, d_glo=S * S, tri=3 if masked else 0,
                        accumulate=True, dtype_in=dt, backend=be)
            # dx[b,t,h,k] = sum_{s>=t} M[h,s,t] dy[b,s,h,k]
            dx = torch.empty_like(xin)
            native.gemm(native.operand(wt, S, 1, s_glo=S * S), native.operand(dy, 1, F, s_glo=K, s_ghi=S * F), dx, S, K,
                        S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=2 if masked else 0, dtype_in=dt,
                        backend=be)
            return dx

        params.tape.append(TapeEntry(bwd))
    return Tensor(out, x.dims)


def _general_attention(args: BlockArgs) -> Tensor:
    """The full form of src/model/spatial.py:42-81 (see include/obst_b200.h, obst_attn_softmax_fwd, for the soft-max
    stage).  The contractions are obst_gemm calls on the [batch, seq, heads, *] layouts as they lie:
        logits[b,s,h,t] = sum_k q[b,s,h,k] key[b,t,h,k]        out[b,s,h,k] = sum_t p[b,s,h,t] val[b,t,h,k]
    The projections (activated_linear_in / _out, one weight each for key, query and value) record their own tapes;
    one composite tape entry differentiates the whole layer."""
    from .backend import run_tape
    from .basic import activated_linear_in, activated_linear_out
    params = args.params
    x = args.tensor
    dim = get_attention_dim(args).dim
    tmp = anonymize_dim(dim)
    if len(x.dims) != 4 or x.dims[1] != dim or x.dims[2:] != params.feature_dims:
        raise NotImplementedError(f"attention: expected [batch, {dim.name}, heads, features_per_head], got {x.dims}")
    Bn, S, H, K = (d.size for d in x.dims)
    F = H * K
    be = params.gemm_backend
    dt = x.data.dtype
    outer = params.tape
    pending, params.residual = params.residual, None
    masked = is_masked(args)

    def branch(fn):
        params.tape = [] if outer is not None else None
        try:
            y = fn()
        finally:
            tape, params.tape = params.tape, outer
        return y, tape

    def masked_map():
        var = embed(args, [params.head_dim, dim, tmp])
        var.set_masked(dim.size if masked else 0)
        return var

    base = tape_in = None
    if "dot_product" in args or "input_as_value" not in args:
        base, tape_in = branch(lambda: activated_linear_in(args))
    bargs = args(base) if base is not None else None

    qry = key = pos = None
    tape_q = tape_k = None
    val = tape_v = None
    logits = None
    scale = 1.0
    if 'dot_product' in args:
        if 'embedded' in args or 'context' in args:
            key, tape_k = branch(lambda: activated_linear_out(bargs))
        if 'embedded' in args or 'positional' in args:
            pos = embed(args, [dim] + params.feature_dims)
        qry, tape_q = branch(lambda: activated_linear_out(bargs))
        scale = float(dim.size) ** -0.5  # qry *= dim.size ** -0.5 (:61), applied inside the soft-max kernel
        if key is not None and pos is not None:
            summed = torch.empty_like(key.data)
            native.bias_add(key.data, pos.value, summed)
            key_data = summed
        elif key is not None:
            key_data = key.data
        elif pos is not None:
            key_data = pos.value  # [seq, heads, features]: the same keys for every batch row
        else:
            raise ValueError("dot_product attention needs 'embedded', 'context' or 'positional' keys")
        key_batched = key is not None
        logits = torch.empty((Bn, S, H, S), dtype=dt, device=x.data.device)
        native.gemm(native.operand(qry.data, F, 1, s_glo=K, s_ghi=S * F),
                    native.operand(key_data, F, 1, s_glo=K, s_ghi=S * F if key_batched else 0), logits, S, S, K, GLO=H,
                    GHI=Bn, d_ld=H * S, d_glo=S, d_ghi=S * H * S, dtype_in=dt, backend=be)
        if "shared_key_value" in args:
            val = key_data
    bias_var = masked_map() if 'biased_softmax' in args else None
    if logits is None and bias_var is not None:
        logits = torch.zeros((Bn, S, H, S), dtype=dt, device=x.data.device)
    if logits is None:
        raise NotImplementedError(f"attention without soft-max logits and with {args.name_extras}: only the plain "
                                  f"'biased_attention_map' + 'input_as_value' form is supported")
    add_var = masked_map() if 'biased_attention_map' in args else None
    mul_var = masked_map() if 'scale_attention_map' in args else None
    soft = torch.empty_like(logits)
    probs = torch.empty_like(logits) if (add_var is not None or mul_var is not None) else None
    native.attn_softmax_fwd(logits, None if bias_var is None else bias_var.value,
                            None if add_var is None else add_var.value, None if mul_var is None else mul_var.value,
                            soft, probs, Bn, S, H, S, scale, masked)
    del logits
    p = probs if probs is not None else soft
    if val is None:
        if "input_as_value" in args:
            val = x.data
        else:
            v, tape_v = branch(lambda: activated_linear_out(bargs))
            val = v.data
    params.residual = None
    # `shared_key_value` (src/model/spatial.py:63-64,76-81): `val = key` is set BEFORE the `if val == 0` that would
    # rename its sequence dim, so the final einsum([map, val], shape) finds `sequence` in both operands and in the
    # output -- it only contracts the map's own `_sequence`: out[b,s,h,:] = (sum_t map[b,s,h,t]) * key[b,s,h,:].
    # No token is mixed (a pure soft-max map sums to 1: out = key).  That is what the reference computes, so that is
    # what runs here.
    degenerate = "shared_key_value" in args and 'dot_product' in args
    rowsum_b = None
    if degenerate:
        rows = Bn * S * H
        rowsum = torch.empty((rows,), dtype=dt, device=x.data.device)
        native.sum_heads(p, rowsum, rows, S, 1)                       # sum over t
        rowsum_b = torch.empty((Bn, S, H, K), dtype=dt, device=x.data.device)
        native.broadcast_heads(rowsum, rowsum_b, rows, K, 1)          # ... spread over the features
        if not key_batched:
            val = val.unsqueeze(0).expand(Bn, S, H, K).contiguous()
        f = torch.empty_like(x.data)
        native.mul(rowsum_b, val, f)
        if pending is not None:
            buf, sign, out = pending
            native.axpby(buf, f, sign, out)
        else:
            out = f
    else:
        p_op = native.operand(p, H * S, 1, s_glo=S, s_ghi=S * H * S)
        v_op = native.operand(val, 1, F, s_glo=K, s_ghi=S * F)
        if pending is not None:
            buf, sign, out = pending
            native.gemm(p_op, v_op, out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, R=buf,
                        epi=2 if sign > 0 else 3, dtype_in=dt, backend=be)
        else:
            out = torch.empty_like(x.data)
            native.gemm(p_op, v_op, out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, dtype_in=dt, backend=be)

    if outer is not None:
        def into(tape, g):
            # a branch that ends in a ReLU fused into its GEMM leaves the [y > 0] mask to whoever produces dy
            if tape and tape[-1].out_relu is not None:
                m = torch.empty_like(g)
                native.act_bwd(0, tape[-1].out_relu, g, m)
                return m
            return g

        def bwd(dy, mask, acc):
            dp = torch.empty_like(soft)
            dval = torch.empty_like(val)
            if degenerate:  # out = rowsum(map) * key
                native.mul(rowsum_b, dy, dval)
                tmp = torch.empty_like(val)
                native.mul(dy, val, tmp)
                d_rowsum = torch.empty((Bn * S * H,), dtype=dt, device=dy.device)
                native.sum_heads(tmp, d_rowsum, Bn * S * H, K, 1)
                native.broadcast_heads(d_rowsum, dp, Bn * S * H, S, 1)
            else:  # through out = p @ val
                native.gemm(native.operand(dy, F, 1, s_glo=K, s_ghi=S * F), native.operand(val, F, 1, s_glo=K, s_ghi=S * F),
                            dp, S, S, K, GLO=H, GHI=Bn, d_ld=H * S, d_glo=S, d_ghi=S * H * S, dtype_in=dt, backend=be)
                native.gemm(native.operand(p, 1, H * S, s_glo=S, s_ghi=S * H * S),
                            native.operand(dy, 1, F, s_glo=K, s_ghi=S * F), dval, S, K, S, GLO=H, GHI=Bn, d_ld=F,
                            d_glo=K, d_ghi=S * F, dtype_in=dt, backend=be)
            # through the soft-max stage: dp -> dlogits (in place), map gradients
            native.attn_softmax_bwd(dp, soft, None if add_var is None else add_var.value,
                                    None if mul_var is None else mul_var.value, dp,
                                    None if bias_var is None else bias_var.ensure_grad(),
                                    None if add_var is None else add_var.ensure_grad(),
                                    None if mul_var is None else mul_var.ensure_grad(), Bn, S, H, S, scale, masked)
            dbase = None
            dx = None
            if qry is not None:
                dq = torch.empty_like(qry.data)
                native.gemm(native.operand(dp, H * S, 1, s_glo=S, s_ghi=S * H * S),
                            native.operand(key_data, 1, F, s_glo=K, s_ghi=S * F if key_batched else 0), dq, S, K, S,
                            GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, dtype_in=dt, backend=be)
                dk = torch.empty_like(qry.data)
                native.gemm(native.operand(dp, 1, H * S, s_glo=S, s_ghi=S * H * S),
                            native.operand(qry.data, 1, F, s_glo=K, s_ghi=S * F), dk, S, K, S, GLO=H, GHI=Bn, d_ld=F,
                            d_glo=K, d_ghi=S * F, dtype_in=dt, backend=be)
                if "shared_key_value" in args:
                    native.axpby(dk, dval, 1.0, dk)
                if pos is not None and pos.ensure_grad() is not None:
                    native.colsum_accumulate(dk, pos.ensure_grad())
                dbase = run_tape(tape_q, into(tape_q, dq))
                if tape_k is not None:
                    dbase = run_tape(tape_k, into(tape_k, dk), acc=dbase)
            if tape_v is not None:
                g = into(tape_v, dval)
                dbase = run_tape(tape_v, g, acc=dbase) if dbase is not None else run_tape(tape_v, g)
            elif "input_as_value" in args and "shared_key_value" not in args:
                dx = dval
            if dbase is not None:
                via_base = run_tape(tape_in, into(tape_in, dbase), acc=dx)
                dx = via_base
            return dx

        outer.append(TapeEntry(bwd))
    return Tensor(out, x.dims)


def _running_sum(args: BlockArgs, mean: bool) -> Tensor:
    """cumsum / cummean along the attention dim (src/model/spatial.py:26-39): out[s] = sum_{t<=s} x[t] (divided by
    s + 1 for cummean); gradient = the reverse running sum of dy (of dy / (t + 1) for cummean)."""
    x = args.tensor
    dim = get_attention_dim(args).dim
    idx = x.dims.index(dim)
    outer = 1
    for d in x.dims[:idx]:
        outer *= d.size
    inner = 1
    for d in x.dims[idx + 1:]:
        inner *= d.size
    seq = dim.size
    xin = x.data.contiguous()
    out = torch.empty_like(xin)
    native.cumsum(xin, out, outer, seq, inner, reverse=False, scale=1 if mean else 0)
    if args.params.tape is not None:
        def bwd(dy, mask, acc):
            dx = torch.empty_like(xin)
            native.cumsum(dy.contiguous(), dx, outer, seq, inner, reverse=True, scale=2 if mean else 0)
            return dx

        args.params.tape.append(TapeEntry(bwd))
    return Tensor(out, x.dims)


def cumsum(args: BlockArgs) -> Tensor:
    return _running_sum(args, False)


def cummean(args: BlockArgs) -> Tensor:
    return _running_sum(args, True)
