"""Model assembly (src/model/__init__.py of the reference): build = _input -> _body -> _output -> _loss.

Language ("gpt") branch only -- video and contrastive branches are outside the hot path.  Everything executes
immediately on the current CUDA stream; when `params.tape` is a list, each executed layer appends its backward.
"""
import typing

import torch

from .. import native
from ..dataclass import BlockArgs, ModelParameter
from ..utils_core import scoped
from ..utils_mtf import Dimension, Tensor, shape_size
from .backend import TapeEntry, linear_to_features, run_tape
from .embedding import embed, gather_embed
from .frontend import LAYER_FUNCTIONS, block_part_fn  # noqa: F401
from .revnet import checkpoint_body, plain_body, reversible_body


class Loss:
    """Scalar results of one build(): device-side partial sums plus the normaliser; `.item()`-style access
    happens once per step on the host (src/run/train.py:14-16 casts the loss to float64)."""

    def __init__(self, sums: torch.Tensor, n_global: int, z_loss: float):
        self.sums = sums  # double[3]: sum nll, sum logZ^2, #correct -- of THIS shard / micro-batch
        self.n_global = n_global
        self.z_loss = z_loss

    @staticmethod
    def values(sums_host, n_global: int, z_loss: float) -> typing.Tuple[float, float, float]:
        nll = float(sums_host[0]) / n_global
        loss = nll + z_loss * float(sums_host[1]) / n_global
        return loss, nll, float(sums_host[2]) / n_global


def _input(params: ModelParameter, txt_src: Tensor) -> Tensor:
    """src/model/__init__.py:75-85: token gather-embed (factorised vocab) + linear to the feature dims."""
    if params.use_video:
        raise NotImplementedError("video input is outside the hot path (model_mode 'gpt' disables it)")
    base_args = BlockArgs(params, txt_src, [''])
    intermediate = params.intermediate[0]
    intermediate = Dimension(intermediate.name, int(intermediate.size * params.vocab_weight_factorization))
    txt = gather_embed(base_args(txt_src, params.token_embedding), [params.vocab_dim, intermediate])
    if params.input_dropout:
        raise NotImplementedError("input_dropout > 0 is not on the hot path")
    txt = linear_to_features(base_args(txt), [params.token_patch_dim, intermediate])
    for config_idx, config in enumerate(params.input_block_config):
        txt = block_part_fn(params, config, txt, f'lang_inp{config_idx}')
    return txt


def _initial_position_embedding(params: ModelParameter, src: Tensor) -> Tensor:
    """src += embed(position_embedding, [dim] + feature_dims) for every position dim after the batch
    (src/model/__init__.py:97-99); text has one, the sequence."""
    position_dims = [d for d in src.dims if d not in params.feature_dims][1:]
    if position_dims != [src.dims[1]] or src.dims[2:] != params.feature_dims:
        raise NotImplementedError(f"initial position embedding over {position_dims} of {src.dims}")
    pos = embed(BlockArgs(params, src, [''])(params.position_embedding), position_dims + params.feature_dims)
    out = torch.empty_like(src.data)
    native.bias_add(src.data, pos.value, out)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            grad = pos.ensure_grad()  # None for the sinusoidal constant
            if grad is not None:
                native.colsum_accumulate(dy, grad)
            return dy

        params.tape.append(TapeEntry(bwd))
    return Tensor(out, src.dims)


def _body(params: ModelParameter, src: Tensor) -> Tensor:
    """src/model/__init__.py:94-130"""
    # [batch, sequence, language_token_patch(=1), heads, features] -> [batch, sequence, heads, features]
    if params.token_patch_dim in src.dims:
        if params.token_patch_dim.size != 1:
            raise NotImplementedError("token_patch_size > 1 is not on the hot path")
        dims = [d for d in src.dims if d != params.token_patch_dim]
        src = Tensor(src.data.view([d.size for d in dims]), dims)
    if params.use_initial_position_embedding:  # src/model/__init__.py:97-99
        src = _initial_position_embedding(params, src)
    strategy = params.memory_reduction_strategy
    if strategy == "revnet":
        return reversible_body(params, src)
    if strategy == "checkpoint":
        return checkpoint_body(params, src)
    if strategy == "none":
        return plain_body(params, src)
    raise NotImplementedError(f"memory_reduction_strategy '{strategy}': 'revnet' (every mixer config), 'checkpoint' and "
                              f"'none' have a B200 path; 'momentum' (MomentumNet) does not")


class FusedLoss:
    """The loss taken in the LM head's own epilogue (obst_gemm epilogue 12): what `_output` needs to know of it."""

    def __init__(self, targets: torch.Tensor, sums: torch.Tensor, n_global: int, z_loss: float):
        self.targets, self.sums, self.n_global, self.z_loss = targets, sums, n_global, z_loss


def _loss_sums(params: ModelParameter, device) -> torch.Tensor:
    sums = getattr(params, "loss_sums", None)  # one accumulator per step when driven by get_train_model
    if sums is None:
        sums = torch.zeros(3, dtype=torch.float64, device=device)
    return sums


def _fused_loss(params: ModelParameter, out: Tensor, txt_tgt: Tensor) -> typing.Optional[FusedLoss]:
    """LM head -> softmax cross-entropy as ONE kernel (src/model/__init__.py:144-145 -> src/mtf_wrapper.py:64-71) when
    the vocabulary is one 256-column tile of the tcgen05 GEMM and the step wants the gradient, not the logits."""
    vocab = params.token_patch_dim.size * params.vocab_dim.size
    tokens = out.data.numel() // shape_size(params.feature_dims)
    if (not getattr(params, "fuse_lm_head_loss", True) or params.tape is None or getattr(params, "keep_logits", False)
            or out.data.dtype != torch.bfloat16 or params.gemm_backend == 1 or vocab != 256 or tokens <= 128
            or txt_tgt.data.numel() != tokens or params.contrastive_across_samples
            or params.contrastive_across_token_embeddings):
        return None
    return FusedLoss(txt_tgt.data, _loss_sums(params, out.data.device), params.n_global_targets or txt_tgt.size,
                     float(params.z_loss or 0.0))


def _output(params: ModelParameter, out: Tensor, fused: typing.Optional[FusedLoss] = None) -> Tensor:
    """src/model/__init__.py:138-145: logits = einsum(out[b,s,h,k], W_o[h,k,p,v]).  With `fused`, the tensor returned
    holds the loss's gradient with respect to the logits instead (they are never written)."""
    base_args = BlockArgs(params, out, [''])
    for config_idx, config in enumerate(params.output_block_config):
        out = block_part_fn(params, config, out, f'lang_out{config_idx}')
    new = [params.token_patch_dim, params.vocab_dim]
    old = params.feature_dims
    var = embed(base_args(params.output_embedding), old + new)
    x = out.data
    F, V = shape_size(old), shape_size(new)
    T = x.numel() // F
    logits = torch.empty(list(x.shape[:-2]) + [d.size for d in new], dtype=x.dtype, device=x.device)
    be, dt, w = params.gemm_backend, x.dtype, var.value
    if fused is not None:
        native.gemm(native.operand(x, F, 1), native.operand(w, 1, V), logits, T, V, F, d_ld=V, dtype_in=dt, backend=be,
                    epi=12, aux=(fused.targets, None, None, None, fused.sums),
                    scalars=(1.0 / fused.n_global, fused.z_loss))
    else:
        native.gemm(native.operand(x, F, 1), native.operand(w, 1, V), logits, T, V, F, d_ld=V, dtype_in=dt, backend=be)
    if params.tape is not None:
        def bwd(dy, mask, acc):
            native.gemm(native.operand(x, 1, F), native.operand(dy, 1, V), var.ensure_grad(), F, V, T, d_ld=V,
                        accumulate=True, dtype_in=dt, backend=be)
            dx = torch.empty_like(x)
            native.gemm(native.operand(dy, V, 1), native.operand(w, V, 1), dx, T, F, V, d_ld=F, dtype_in=dt,
                        backend=be)
            return dx

        params.tape.append(TapeEntry(bwd))
    return Tensor(logits, [d for d in out.dims if d not in old] + new)


def _loss(params: ModelParameter, token_out: Tensor, txt_tgt: Tensor, fused: typing.Optional[FusedLoss] = None) -> Loss:
    """softmax cross-entropy + z-loss + accuracy (src/mtf_wrapper.py:64-71, src/model/__init__.py:183-187).
    The normaliser is targets.size of the GLOBAL batch (all ranks, all micro-batches)."""
    if params.contrastive_across_samples or params.contrastive_across_token_embeddings:
        raise NotImplementedError("contrastive losses are outside the hot path")
    if fused is not None:  # the LM head's epilogue has done it: token_out IS the gradient
        dlogits = token_out.data
        params.tape.append(TapeEntry(lambda dy, mask, acc: dlogits))
        return Loss(fused.sums, fused.n_global, fused.z_loss)
    n_global = params.n_global_targets or txt_tgt.size
    logits = token_out.data
    sums = _loss_sums(params, logits.device)
    if params.tape is not None and getattr(params, "keep_logits", False):
        # the gradient overwrites the logits in place below; a caller that wants `token_out` back from a training step
        # (the reference returns it, src/model/__init__.py:229) asks for this copy -- training loops do not
        token_out.data = logits.clone()
    dlogits = logits  # gradient overwrites the logits in place
    native.xent_fwd_bwd(logits, txt_tgt.data, dlogits, sums, 1.0 / n_global, float(params.z_loss or 0.0))
    if params.tape is not None:
        params.tape.append(TapeEntry(lambda dy, mask, acc: dlogits))
    return Loss(sums, n_global, float(params.z_loss or 0.0))


def _build(params: ModelParameter, txt_src: Tensor, txt_tgt: Tensor):
    src = scoped("input", _input, params, txt_src)
    out = scoped("body", _body, params, src)
    if getattr(params, "skip_loss", False):  # sampling (obst.run.inference): the logits themselves are the result
        token_out = scoped("output", _output, params, out)
        params.attention_idx = 0
        return None, token_out
    fused = _fused_loss(params, out, txt_tgt)
    token_out = scoped("output", _output, params, out, fused)
    token_loss = scoped("loss", _loss, params, token_out, txt_tgt, fused)
    params.attention_idx = 0
    return token_loss, token_out


def build(params: ModelParameter,
          vid: typing.Optional[Tensor],
          cat_msk_src: typing.Optional[Tensor],
          cat_msk_tgt: typing.Optional[Tensor],
          txt_src: typing.Optional[Tensor],
          txt_tgt: typing.Optional[Tensor],
          vid_msk_src: typing.Optional[Tensor],
          vid_msk_tgt: typing.Optional[Tensor],
          txt_msk: typing.Optional[Tensor]):
    """Same signature and return tuple as the reference's build (src/model/__init__.py:231-259):
    (loss, loss_list, video_loss, accuracy, token_loss, frame_out, token_out).  `loss`/`token_loss` are `Loss`
    objects (device sums + normaliser); `accuracy` is read from the same object when calc_accuracy is set."""
    if vid is not None or params.use_video:
        raise NotImplementedError("video inputs are outside the hot path; use model_mode='gpt' (src/main.py:87-93)")
    if not params.use_language or txt_src is None or txt_tgt is None:
        raise ValueError("the language path needs txt_src and txt_tgt")
    token_loss, token_out = scoped(params.model_mode, _build, params, txt_src, txt_tgt)
    accuracy = token_loss if params.calc_accuracy else None
    return token_loss, [token_loss], None, accuracy, token_loss, None, token_out


def backward(params: ModelParameter):
    """Reverse sweep over the recorded tape (the part of src/optimizer/__init__.py:128-174 that differentiates the
    graph).  Parameter gradients are accumulated into Variable.grad (fp32)."""
    tape, params.tape = params.tape, None
    if not tape:
        raise RuntimeError("backward() without a recorded forward pass (params.tape was None)")
    run_tape(tape, None)
