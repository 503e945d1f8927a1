"""norm(): src/model/normalization.py:22-34 -- mean-subtract, rsqrt(mean-square + 1e-5), optional learned scale
(N(1,0.02)) and shift (N(0,0.02)); with `group` each head is normalised on its own.  One fused kernel forward,
one backward (obst_gn_fwd / obst_gn_bwd)."""
import typing

import torch

from .. import native
from ..dataclass import BlockArgs
from ..utils_mtf import DIM_LIST, Tensor, linear_shapes, shape_size, shape_sub
from .backend import TapeEntry, normal_var


def norm_layout(args: BlockArgs, feature_shape: typing.Optional[DIM_LIST] = None):
    """What norm() normalises over, and its scale / shift variables (created here, in the reference's order):
    returns (heads, width, gamma, beta)."""
    params = args.params
    x = args.tensor
    feature_shape = list(linear_shapes(args).old if feature_shape is None else feature_shape)
    reduced = shape_sub(feature_shape, [params.head_dim] * ('group' in args))
    # dims the statistics are taken over must be the trailing (contiguous) dims; the rest of feature_shape
    # (the head dim under `group`) must sit directly in front of them
    if not reduced or x.dims[-len(reduced):] != reduced:
        raise NotImplementedError(f"norm: normalised dims {reduced} must be the trailing dims of {x.dims}")
    outer = shape_sub(feature_shape, reduced)
    if outer and x.dims[-len(reduced) - len(outer):-len(reduced)] != outer:
        raise NotImplementedError(f"norm: {outer} must directly precede {reduced} in {x.dims}")
    width = shape_size(reduced)
    heads = shape_size(outer)
    gamma = normal_var(args, feature_shape, mean=1) if 'scale' in args else None
    beta = normal_var(args, feature_shape, mean=0) if 'shift' in args else None
    return heads, width, gamma, beta


def norm(args: BlockArgs, feature_shape: typing.Optional[DIM_LIST] = None, fuse_gelu: bool = False) -> Tensor:
    params = args.params
    x = args.tensor
    heads, width, gamma, beta = norm_layout(args, feature_shape)
    act = 1 if fuse_gelu else 0
    xin = x.data
    y = torch.empty_like(xin)
    # (mean, rstd) per normalised row, kept for the backward kernel when one will run (8 B per `width` elements)
    stats = None
    if params.tape is not None:
        stats = torch.empty(xin.numel() // width, 2, dtype=torch.float32, device=xin.device)
    native.gn_fwd(xin, None if gamma is None else gamma.value, None if beta is None else beta.value, y, heads, width,
                  act, stats=stats)
    if params.tape is not None:
        def bwd(dy: torch.Tensor, mask, acc):
            dx = acc if acc is not None else torch.empty_like(xin)
            native.gn_bwd(dy, xin, None if gamma is None else gamma.value, None if beta is None else beta.value, dx,
                          None if gamma is None else gamma.ensure_grad(),
                          None if beta is None else beta.ensure_grad(), heads, width, act=act,
                          relu_mask=mask is not None, dx_accumulate=acc is not None, stats=stats)
            return dx

        # a ReLU mask can only be folded when the masking tensor IS this layer's input (norm right after relu)
        params.tape.append(TapeEntry(_mask_guard(bwd, xin), supports_mask=True, supports_acc=True))
    return Tensor(y, x.dims)


def _mask_guard(bwd, xin):
    def fn(dy, mask, acc):
        if mask is not None and mask.data_ptr() != xin.data_ptr():
            dx = bwd(dy, None, None)
            native.act_bwd(0, mask, dx, dx)
            if acc is not None:
                native.axpby(acc, dx, 1.0, acc)
                return acc
            return dx
        return bwd(dy, mask, acc)

    return fn
