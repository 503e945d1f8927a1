"""Reversible residual stack: the RevGradOp of src/model/revnet.py, run eagerly.

forward  (src/model/revnet.py:20-24):  (x1, x2) -> (x2, x1 + f(x2)); nothing but the two streams is kept.
backward (src/model/revnet.py:55-120): for each sub-block, last to first: x2 is known, f(x2) is RE-COMPUTED with
the tape on, x1 = y1 - f(x2) is reconstructed (the subtraction rides on the last GEMM's epilogue), then f is
back-propagated: dx1 = dy1, dx2 = dy2 + J^T dy1 (the add rides on the first norm's backward kernel).
All four buffers (two streams, two gradients) are updated in place."""
import typing

import torch

from .. import native
from .. import utils_core
from ..dataclass import BlockConfig, ModelParameter
from ..utils_mtf import Tensor
from .backend import TapeEntry, reset_shared_ordinals, run_tape
from .frontend import block_part_fn


def _apply(params: ModelParameter, config: BlockConfig, index: int, x2: Tensor, x1_buf: torch.Tensor, sign: float,
           out: typing.Optional[torch.Tensor] = None):
    """out <- x1_buf + sign * f(x2) (out defaults to x1_buf: in place), fused into f's last GEMM when that layer
    supports it."""
    out = x1_buf if out is None else out
    params.residual = (x1_buf, sign, out)
    f = block_part_fn(params, config, x2, str(index))
    if params.residual is not None:  # last layer could not fuse the residual
        params.residual = None
        native.axpby(x1_buf, f.data, sign, out)


def stored_body(params: ModelParameter, src: Tensor) -> Tensor:
    """EXTENSION (`revnet_store_activations: true`, not a reference mode): the same reversible wiring
    (x1, x2) -> (x2, x1 + f(x2)), but every sub-block's intermediates are KEPT (180 GB of HBM hold ~40 k tokens x
    32 depths), so backward neither recomputes f nor reconstructs x1.  Gradients differ from the reversible path
    only by the reconstruction rounding the reference incurs (they are closer to the exact ones)."""
    dims = src.dims
    p, q = src.data, src.data
    outer_tape = params.tape
    tapes: typing.List[list] = []
    for i in range(params.depth):
        for config in params.block_config:
            sub: list = [] if outer_tape is not None else None
            params.tape = sub
            y = torch.empty_like(p)
            _apply(params, config, i, Tensor(q, dims), p, 1.0, out=y)
            tapes.append(sub)
            p, q = q, y
    out = torch.empty_like(p)
    native.axpby(p, q, 1.0, out)
    params.tape = outer_tape
    if outer_tape is not None:
        n_cfg = len(params.block_config)

        def bwd(dout: torch.Tensor, mask, acc):
            dp, dq = dout, dout.clone()
            reducer = getattr(params, "grad_reducer", None)
            for j in range(len(tapes) - 1, -1, -1):
                run_tape(tapes[j], dq, acc=dp)  # dp <- dy2 + J^T dy1
                tapes[j] = None                 # frees this sub-block's activations
                dp, dq = dq, dp
                if j % n_cfg == 0 and reducer is not None:
                    reducer.on_depth_done(j // n_cfg)
            native.axpby(dp, dq, 1.0, dp)
            return dp

        outer_tape.append(TapeEntry(bwd))
    return Tensor(out, dims)


def reversible_body(params: ModelParameter, src: Tensor) -> Tensor:
    """_body for memory_reduction_strategy == 'revnet' (src/model/__init__.py:101-106,124-129)."""
    if getattr(params, "revnet_store_activations", False):
        return stored_body(params, src)
    dims = src.dims
    p = src.data.clone()  # x1
    q = src.data          # x2 (src itself is not needed again: its gradient path is dp + dq)
    outer_tape = params.tape
    snapshots: typing.List[typing.Tuple[dict, int]] = []
    params.tape = None  # forward keeps no activations
    trace = getattr(params, "debug_trace", None)  # tests only: copies of the two streams after every sub-block
    if trace is not None:
        trace["src"] = src.data.clone()
    for i in range(params.depth):
        for config in params.block_config:
            snapshots.append((utils_core.snapshot(), params.attention_idx))
            _apply(params, config, i, Tensor(q, dims), p, 1.0)  # y1 = x1 + f(x2), in place on x1's buffer
            p, q = q, p                                          # outputs (x2, y1)
            if trace is not None:
                trace[f"x1_{i}_{len(snapshots) - 1 - i * len(params.block_config)}"] = p.clone()
                trace[f"x2_{i}_{len(snapshots) - 1 - i * len(params.block_config)}"] = q.clone()
    out = torch.empty_like(p)
    native.axpby(p, q, 1.0, out)  # out = x1 + x2  (src/model/__init__.py:129)
    if trace is not None:
        trace["out"] = out.clone()
    params.tape = outer_tape

    if outer_tape is not None:
        n_cfg = len(params.block_config)

        def bwd(dout: torch.Tensor, mask, acc):
            nonlocal p, q
            dp = dout
            dq = dout.clone()
            saved_names, saved_idx = utils_core.snapshot(), params.attention_idx
            saved_scope = list(utils_core._SCOPE)
            reducer = getattr(params, "grad_reducer", None)
            utils_core._SCOPE[:] = body_scope
            for j in range(len(snapshots) - 1, -1, -1):
                i, c = divmod(j, n_cfg)
                config = params.block_config[c]
                names, att_idx = snapshots[j]
                utils_core.restore(names)
                params.attention_idx = att_idx
                reset_shared_ordinals(params, str(i), f"{c}v")
                # state after block j is (p, q) = (x2_in, x1_in + f(x2_in))
                params.tape = []
                params.upstream_grad = dq  # the gradient f's output will receive: layers may fold reductions with it
                try:
                    _apply(params, config, i, Tensor(p, dims), q, -1.0)   # q <- x1_in = y1 - f(x2_in)
                finally:
                    params.upstream_grad = None
                tape, params.tape = params.tape, None
                run_tape(tape, dq, acc=dp)                              # dp <- dy2 + J^T dy1
                del tape
                p, q, dp, dq = q, p, dq, dp                             # (x1_in, x2_in), (dx1, dx2)
                if trace is not None:
                    trace[f"rx1_{i}_{c}"] = p.clone()
                    trace[f"rx2_{i}_{c}"] = q.clone()
                if c == 0 and reducer is not None:
                    reducer.on_depth_done(i)  # both sub-blocks of depth i done: its gradients are final
            utils_core.restore(saved_names)
            params.attention_idx = saved_idx
            utils_core._SCOPE[:] = saved_scope
            native.axpby(dp, dq, 1.0, dp)  # d src = dx1 + dx2 (x1 = x2 = src at the input)
            return dp

        body_scope = list(utils_core._SCOPE)
        outer_tape.append(TapeEntry(bwd))
    return Tensor(out, dims)


def plain_body(params: ModelParameter, src: Tensor) -> Tensor:
    """_body for memory_reduction_strategy == 'none' (src/model/__init__.py:119-123): the sub-blocks are simply
    composed, out = f_n(... f_1(src)), every layer's backward stays on the tape (a sub-block adds its input when its
    block_config sets `skip`, src/model/frontend.py:29-30)."""
    out = src
    for i in range(params.depth):
        for config in params.block_config:
            out = block_part_fn(params, config, out, str(i))
    return out


def checkpoint_body(params: ModelParameter, src: Tensor) -> Tensor:
    """_body for memory_reduction_strategy == 'checkpoint' (src/model/__init__.py:107-112, mtf recompute_grad): the
    same composition as 'none', but forward keeps only each sub-block's INPUT; backward re-runs the sub-block from it
    with the tape on, then back-propagates -- the recompute of the reversible path without its reconstruction."""
    dims = src.dims
    outer_tape = params.tape
    params.tape = None
    x = src.data
    records: typing.List[typing.Tuple[dict, int, torch.Tensor]] = []
    for i in range(params.depth):
        for config in params.block_config:
            records.append((utils_core.snapshot(), params.attention_idx, x))
            # mtf_wrapper.recompute_grad is itself `scoped`: every block sits one scope level deeper than under the
            # other strategies ("body0v/recompute_grad<n>v/<depth>_<config>v/..."), and that is part of the names
            x = utils_core.scoped("recompute_grad", block_part_fn, params, config, Tensor(x, dims), str(i)).data
    params.tape = outer_tape
    if outer_tape is not None:
        n_cfg = len(params.block_config)
        body_scope = list(utils_core._SCOPE)

        def bwd(dy: torch.Tensor, mask, acc):
            saved_names, saved_idx = utils_core.snapshot(), params.attention_idx
            saved_scope = list(utils_core._SCOPE)
            reducer = getattr(params, "grad_reducer", None)
            utils_core._SCOPE[:] = body_scope
            for j in range(len(records) - 1, -1, -1):
                i, c = divmod(j, n_cfg)
                names, att_idx, xin = records[j]
                utils_core.restore(names)
                params.attention_idx = att_idx
                reset_shared_ordinals(params, str(i), f"{c}v")
                params.tape = []
                utils_core.scoped("recompute_grad", block_part_fn, params, params.block_config[c], Tensor(xin, dims), str(i))
                tape, params.tape = params.tape, None
                dy = run_tape(tape, dy)
                records[j] = None  # frees the stored input
                if c == 0 and reducer is not None:
                    reducer.on_depth_done(i)
            utils_core.restore(saved_names)
            params.attention_idx = saved_idx
            utils_core._SCOPE[:] = saved_scope
            return dy

        outer_tape.append(TapeEntry(bwd))
    return Tensor(x, dims)
