"""Training-step driver: get_train_model of src/run/train.py:19-77, executed instead of graph-built.

The caller contract is the reference's: token inputs `token_x`/`token_y` of shape [batch, sequence,
language_token_patch] (x = window[:, :S], y = window[:, 1:S+1], src/inputs.py:543-551), one call = one optimizer
step, returns the same tuple.  The rows given here are THIS rank's shard of the global batch; they are processed
in micro-batches of `params.micro_batch_size` rows (an execution detail: the model is per-token, the loss is
normalised by the global token count, gradients accumulate in fp32), the reference's "add ... add, update"
sequence of get_optimizer calls."""
import typing

import torch

from .. import utils_core
from ..dataclass import ModelParameter
from ..model import Loss, build
from ..optimizer import get_optimizer
from ..utils_mtf import Tensor


def _as_tensor(params: ModelParameter, t) -> typing.Optional[Tensor]:
    if t is None or isinstance(t, Tensor):
        return t
    if t.dim() == 2:
        t = t.unsqueeze(-1)
    dims = [params.batch_dim._replace(size=t.shape[0]), params.sequence_dim, params.token_patch_dim]
    return Tensor(t.contiguous(), dims)


def get_train_model(params: ModelParameter, frame_input, cat_mask_src, cat_mask_tag, token_x_input, token_y_input,
                    frame_mask_src, frame_mask_tag, token_mask, manual_global_step):
    """One optimizer step (src/run/train.py:19-77).

    macro_batching == 1: one build + one get_optimizer(..., "update").
    macro_batching  > 1: the input carries train_batch_size * macro_batching rows; row b * macro + s belongs to slice
    s (reshape to [batch, batch_slice], unbind: `:22-33`).  Every slice is built and differentiated on its own; all
    but the last run get_optimizer(..., "add"): the optimizer state moves on, the master weights do not, and the
    next slice's graph reads var.value - update (`:63-69`); the last runs "update".  The returned loss is the mean
    over the slices, first_loss / last_loss those of the first / last slice (`:44-50,75-77`).

    Independently of that, the rows of one slice are processed in micro-batches of `params.micro_batch_size` rows (an
    execution detail of this implementation: the model is per-token, the loss is normalised by the slice's global
    token count, gradients accumulate in fp32 -- get_optimizer(..., "accumulate"))."""
    if frame_input is not None:
        raise NotImplementedError("video inputs are outside the hot path")
    tx, ty = _as_tensor(params, token_x_input), _as_tensor(params, token_y_input)
    macro = int(params.macro_batching)
    if tx.data.shape[0] % macro:
        raise ValueError(f"{tx.data.shape[0]} input rows cannot be split into {macro} macro-batches")
    world = getattr(getattr(params, "grad_reducer", None), "world", 1)
    dev = tx.data.device
    if getattr(params, "loss_sums", None) is None:
        params.loss_sums = torch.zeros(3, dtype=torch.float64, device=dev)
    update_ops, learning_rate = [], 0.0
    token_out = None
    slice_sums = []
    for s in range(macro):
        sx = tx.data[s::macro] if macro > 1 else tx.data
        sy = ty.data[s::macro] if macro > 1 else ty.data
        if macro > 1:
            sx, sy = sx.contiguous(), sy.contiguous()
        rows = sx.shape[0]
        mb = min(params.micro_batch_size or rows, rows)
        params.n_global_targets = rows * world * params.sequence_length * params.token_patch_size
        params.loss_sums.zero_()
        state = getattr(params, "optimizer_state", None)
        if state is not None:
            state.zero_grad()
        params.macro_chained = s > 0  # this slice's graph reads the weights chained from the previous slice
        starts = list(range(0, rows, mb))
        for i, r0 in enumerate(starts, 1):
            last_micro = i == len(starts)
            sl = slice(r0, min(rows, r0 + mb))
            utils_core.NAME_INDICES.clear()          # src/run/train.py:38-39
            params.cached_parameters.clear()
            params.tape = []
            bx = Tensor(sx[sl], [tx.dims[0]._replace(size=sl.stop - sl.start)] + tx.dims[1:])
            by = Tensor(sy[sl], [ty.dims[0]._replace(size=sl.stop - sl.start)] + ty.dims[1:])
            loss, loss_list, video_loss, accuracy, token_loss, frame_out, token_out = build(
                params, None, None, None, bx, by, None, None, None)
            fn = "accumulate" if not last_micro else ("update" if s == macro - 1 else "add")
            update_ops, learning_rate = get_optimizer(loss_list, params, manual_global_step + s, fn)
        if macro > 1:
            slice_sums.append(params.loss_sums.clone())
    params.macro_chained = False
    params.global_step += params.macro_batching            # src/run/run.py:139-156
    params.manual_global_step += params.macro_batching
    if getattr(params, "restore_entries", None) is not None:  # resumed: every variable and slot exists by now
        from .checkpoint import finish_restore
        finish_restore(params)
    n = params.n_global_targets
    z = float(params.z_loss or 0.0)
    if macro > 1:  # total_loss = mean over the slices (`:50,75`)
        params.loss_sums.copy_(torch.stack(slice_sums).sum(0) / macro)
        first = Loss(slice_sums[0], n, z)
        last = Loss(slice_sums[-1], n, z)
    total = Loss(params.loss_sums, n, z)
    if macro == 1:
        first = last = total
    # token_loss and accuracy are what the LAST macro-batch's build returned (src/run/train.py:40,76-77)
    accuracy = last if params.calc_accuracy else None
    return None, token_out, learning_rate, total, first, last, None, last, accuracy, update_ops, {}


def read_scalars(params: ModelParameter, loss: Loss) -> typing.Dict[str, float]:
    """One device->host read of the step's scalars (the reference writes them as summaries,
    src/run/run.py:123-132): loss, token_loss (nll + z-loss), accuracy."""
    host = loss.sums.cpu()
    value, nll, acc = Loss.values(host, loss.n_global, loss.z_loss)
    return {"loss": value, "token_loss": value, "nll": nll, "accuracy": acc}
