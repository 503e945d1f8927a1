"""Autoregressive sampling: autoregressive_model / get_infrence_model of src/run/inference.py:16-160, language
branch (`params.use_video` is outside the hot path).

The reference wraps `build` in an `mtf.while_loop`: every iteration runs the WHOLE forward pass on the current
token window, perturbs the logits with temperature-scaled Gumbel noise, takes the argmax, shifts it one position to
the right and writes the entry at `position` back into the window (src/run/inference.py:76-95).  Here the loop is a
host loop over the same forward pass (no tape: nothing is kept for a backward), and the noise / argmax / shift /
write-back of one iteration is one kernel (obst_sample_step) that reads the single logits row the reference's
iteration actually uses."""
import typing

import torch

from .. import native, utils_core
from ..dataclass import ModelParameter
from ..model import build
from ..utils_mtf import Tensor
from .train import _as_tensor


def _forward_logits(params: ModelParameter, token_x: Tensor) -> torch.Tensor:
    utils_core.NAME_INDICES.clear()
    params.cached_parameters.clear()
    tape, skip = params.tape, getattr(params, "skip_loss", False)
    params.tape, params.skip_loss = None, True
    try:
        *_, token_out = build(params, None, None, None, token_x, token_x, None, None, None)
    finally:
        params.tape, params.skip_loss = tape, skip
    return token_out.data


def autoregressive_model(params: ModelParameter, frame_input=None, token_x_input=None, token_y_input=None,
                         frame_mask_src=None, frame_mask_tag=None, token_mask=None, initial_pos=None,
                         sampling_temperature=None, end_iterations=None, seed: int = 0, uniform_fn=None
                         ) -> typing.Tuple[torch.Tensor, None]:
    """Same arguments and return pair (token_out, frame_out) as the reference.  `seed` / `uniform_fn` are
    extensions: the reference draws tf.random_uniform, which no other framework reproduces, so the generator here
    is seeded explicitly, and tests may supply the uniform samples of each position (`uniform_fn(position)` ->
    fp32 [batch, vocab])."""
    if params.use_video or frame_input is not None:
        raise NotImplementedError("video sampling is outside the hot path")
    if initial_pos is None:
        initial_pos = params.initial_autoregressive_position
    if sampling_temperature is None:
        sampling_temperature = params.sampling_temperature
    if end_iterations is None:
        end_iterations = params.sequence_length
    tx = _as_tensor(params, token_x_input)
    tokens = tx.data.to(torch.int32).clone().contiguous()
    if params.debug_sample:  # src/run/inference.py:101-107: row 0 as given, row 1 = row 0 masked after initial_pos
        keep = (torch.arange(tokens.shape[1], device=tokens.device) <= int(initial_pos)).to(torch.int32)
        first = tokens[:1]
        tokens = torch.cat([first, first * keep.view(1, -1, *([1] * (tokens.dim() - 2)))], 0).contiguous()
    dims = [tx.dims[0]._replace(size=tokens.shape[0])] + tx.dims[1:]
    position = int(initial_pos)
    while position < int(end_iterations):  # cond_fn, src/run/inference.py:118-122
        logits = _forward_logits(params, Tensor(tokens, dims))
        uniform = None if uniform_fn is None else uniform_fn(position)
        native.sample_step(logits, tokens, position, float(sampling_temperature), seed, uniform)
        position += 1
    return tokens, None


def get_infrence_model(params: ModelParameter):
    """src/run/inference.py:137-160 (the reference's spelling)."""

    def infrence_model(frame_input, cat_mask_src, cat_mask_tag, token_x_input, token_y_input, frame_mask_src,
                       frame_mask_tag, token_mask, initial_pos, sampling_temperature, end_iterations):
        if params.use_autoregressive_sampling:
            return autoregressive_model(params, frame_input, token_x_input, token_y_input, frame_mask_src,
                                        frame_mask_tag, token_mask, initial_pos, sampling_temperature, end_iterations)
        tx = _as_tensor(params, token_x_input)
        return _forward_logits(params, tx), None

    return infrence_model
