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
    if frame_input is not None:
        raise NotImplementedError("video inputs are outside the hot path")
    if int(params.macro_batching) != 1:
        # src/run/train.py:35-74 chains the macro-batches through the weights the previous one WOULD have produced
        # and commits only the last update; that path is not built yet, and silently doing something else (the
        # micro-batches below ACCUMULATE gradients, which is a different algorithm) would be worse than refusing
        raise NotImplementedError(f"macro_batching = {params.macro_batching}: the reference's macro-batching chain "
                                  f"(src/run/train.py:35-74) is not implemented; use macro_batching = 1")
    tx, ty = _as_tensor(params, token_x_input), _as_tensor(params, token_y_input)
    rows = tx.data.shape[0]
    mb = params.micro_batch_size or rows
    mb = min(mb, rows)
    world = getattr(getattr(params, "grad_reducer", None), "world", 1)
    params.n_global_targets = rows * world * params.sequence_length * params.token_patch_size
    if getattr(params, "loss_sums", None) is None:
        params.loss_sums = torch.zeros(3, dtype=torch.float64, device=tx.data.device)
    params.loss_sums.zero_()
    state = getattr(params, "optimizer_state", None)
    if state is not None:
        state.zero_grad()
    starts = list(range(0, rows, mb))
    update_ops, learning_rate = [], 0.0
    token_out = None
    first_sums = None
    for i, r0 in enumerate(starts, 1):
        last = i == len(starts)
        sl = slice(r0, min(rows, r0 + mb))
        utils_core.NAME_INDICES.clear()          # src/run/train.py:38-39
        params.cached_parameters.clear()
        params.tape = []
        bx = Tensor(tx.data[sl], [tx.dims[0]._replace(size=sl.stop - sl.start)] + tx.dims[1:])
        by = Tensor(ty.data[sl], [ty.dims[0]._replace(size=sl.stop - sl.start)] + ty.dims[1:])
        loss, loss_list, video_loss, accuracy, token_loss, frame_out, token_out = build(
            params, None, None, None, bx, by, None, None, None)
        if i == 1:
            first_sums = params.loss_sums.clone() if len(starts) > 1 else None
        update_ops, learning_rate = get_optimizer(loss_list, params, manual_global_step,
                                                  "update" if last else "add")
    params.global_step += params.macro_batching            # src/run/run.py:139-156
    params.manual_global_step += params.macro_batching
    total = Loss(params.loss_sums, params.n_global_targets, float(params.z_loss or 0.0))
    accuracy = total if params.calc_accuracy else None
    return None, token_out, learning_rate, total, first_sums, total, None, total, accuracy, update_ops, {}


def read_scalars(params: ModelParameter, loss: Loss) -> typing.Dict[str, float]:
    """One device->host read of the step's scalars (the reference writes them as summaries,
    src/run/run.py:123-132): loss, token_loss (nll + z-loss), accuracy."""
    host = loss.sums.cpu()
    value, nll, acc = Loss.values(host, loss.n_global, loss.z_loss)
    return {"loss": value, "token_loss": value, "nll": nll, "accuracy": acc}
