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
    params = args.params
    params.attention_idx += 1
    if "dot_product" in args or "input_as_value" not in args or 'biased_softmax' in args \
            or 'scale_attention_map' in args:
        raise NotImplementedError("attention: only 'biased_attention_map' + 'input_as_value' (the mixer path) has a "
                                  "B200 kernel; dot-product / softmax variants are SURVEY.md 8(f) 'next'")
    if 'biased_attention_map' not in args:
        raise UserWarning(f"WARNING: There is no spatial mixing happening with the following attention parameters: "
                          f"{args.name_extras}.")
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
