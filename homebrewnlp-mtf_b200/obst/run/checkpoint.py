"""Checkpoint / resume under the reference's variable names.

The reference relies on TF's Saver over the MTF master variables: every variable is stored under its scope name
(`.../get_variableN`, src/utils_mtf.py:271-280), every optimizer slot under
`<variable>/<optimizer string with ':' -> '_'>/<slot>` (src/optimizer/backend.py:23-25), in the MASTER dtype --
`storage_dtype` for weights (bfloat16 in the 32*_mixer configs, so a restart there resumes from bf16-rounded
weights) and the optimizer's master dtype for slots -- next to the global step.  The TensorBundle file format itself
needs TensorFlow; this bridge keeps the names, dtypes and contents in a single .npz so that a run can be stopped
and resumed (and so that a converter that has TensorFlow only has to rename nothing).

    save_checkpoint(params, path)            after any number of steps
    load_checkpoint(params, path)            on a fresh ModelParameter, BEFORE the first step

`exact=True` stores the fp32 slices instead of the storage dtype: resuming is then bit-identical to not stopping."""
import typing
import warnings

import numpy as np
import torch

from ..dataclass import ModelParameter

_BF16_TAG = "::bf16"


def _pack(t: torch.Tensor, dtype: torch.dtype) -> typing.Tuple[str, np.ndarray]:
    t = t.detach().to(dtype).cpu().contiguous()
    if dtype == torch.bfloat16:
        return _BF16_TAG, t.view(torch.int16).numpy().view(np.uint16)
    return "", t.numpy()


def _unpack(name: str, arr: np.ndarray) -> typing.Tuple[str, torch.Tensor]:
    if name.endswith(_BF16_TAG):
        return name[:-len(_BF16_TAG)], torch.from_numpy(arr.view(np.int16).copy()).view(torch.bfloat16).float()
    return name, torch.from_numpy(np.array(arr))


def save_checkpoint(params: ModelParameter, path: str, exact: bool = False):
    if not params.variable_cache:
        raise RuntimeError("nothing to save: the model has not been built yet")
    out = {"global_step": np.int64(params.global_step), "manual_global_step": np.int64(params.manual_global_step)}
    wdt = torch.float32 if exact else params.variable_dtype.master_dtype
    for name, var in params.variable_cache.items():
        tag, arr = _pack(var.data, wdt)
        out[name + tag] = arr
    state = getattr(params, "optimizer_state", None)
    if state is not None:
        sdt = torch.float32 if exact else params.optimizer_dtype.master_dtype
        for name, slot in state.slots.items():
            tag, arr = _pack(slot, slot.dtype if exact else sdt)
            out[name + tag] = arr
    np.savez(path, **out)


def load_checkpoint(params: ModelParameter, path: str):
    """Arms `params` so that variables (at creation) and optimizer slots (when the optimizer state is created) take
    their values from the file; restores the step counters at once.  A variable or slot the model asks for and the
    file lacks raises KeyError at that point (a file saved before the first update holds no slots at all: those
    start from zero, as in the reference); entries nobody asked for are reported by `finish_restore`, which the
    training step calls once everything exists, and which releases the host copy of the file."""
    with np.load(path if path.endswith(".npz") else path + ".npz") as f:
        entries = dict(_unpack(k, f[k]) for k in f.files)
    params.global_step = int(entries.pop("global_step"))
    params.manual_global_step = int(entries.pop("manual_global_step"))
    params.restore_entries = entries
    params.restore_used = set()
    if params.variable_cache:  # already built: overwrite in place
        for name, var in params.variable_cache.items():
            restore_variable(params, name, var)
        state = getattr(params, "optimizer_state", None)
        if state is not None:
            restore_slots(params, state)
            state.invalidate_weight_norms()
            finish_restore(params)


def take_entry(params: ModelParameter, name: str, kind: str = "variable") -> typing.Optional[torch.Tensor]:
    """The file's tensor for `name`, or None when no restore is armed.  Missing names raise -- except slots of a file
    that was saved before any update (it holds variables only): those keep their zero initialisation."""
    entries = getattr(params, "restore_entries", None)
    if entries is None:
        return None
    if name not in entries:
        if kind == "slot" and not (set(entries) - set(params.variable_cache)):
            return None
        hint = " (optimizer string changed since the file was written?)" if kind == "slot" else ""
        raise KeyError(f"checkpoint has no {kind} '{name}'{hint}")
    params.restore_used.add(name)
    return entries[name]


def finish_restore(params: ModelParameter) -> typing.List[str]:
    """Every variable and slot exists: report file entries that nothing consumed, drop the host copy of the file."""
    entries = getattr(params, "restore_entries", None)
    if entries is None:
        return []
    unused = sorted(set(entries) - params.restore_used)
    params.restore_entries = None
    params.restore_used = set()
    if unused:
        warnings.warn(f"checkpoint entries not used by this model / optimizer: {unused[:8]}"
                      f"{' ...' if len(unused) > 8 else ''} ({len(unused)} in total)")
    return unused


def restore_variable(params: ModelParameter, name: str, var) -> bool:
    src = take_entry(params, name)
    if src is None:
        return False
    if tuple(src.shape) != tuple(var.data.shape):
        raise ValueError(f"checkpoint variable '{name}' has shape {tuple(src.shape)}, model wants {tuple(var.data.shape)}")
    var.data.copy_(src.to(var.data.device, torch.float32))
    var.refresh()
    return True


def restore_slot(params: ModelParameter, name: str, slot: torch.Tensor) -> bool:
    src = take_entry(params, name, "slot")
    if src is None:
        return False
    if src.numel() != slot.numel():
        raise ValueError(f"checkpoint slot '{name}' has {src.numel()} elements, optimizer wants {slot.numel()}")
    slot.copy_(src.to(slot.device).reshape(slot.shape).to(slot.dtype))
    return True


def restore_slots(params: ModelParameter, state):
    if getattr(params, "restore_entries", None) is None:
        return
    for name, slot in state.slots.items():
        restore_slot(params, name, slot)
