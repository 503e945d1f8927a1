"""Checkpoint / resume under the reference's variable names.

The reference relies on TF's Saver over the MTF master variables: every variable is stored under its scope name
(`.../get_variableN`, src/utils_mtf.py:271-280), every optimizer slot under
`<variable>/<optimizer string with ':' -> '_'>/<slot>` (src/optimizer/backend.py:23-25), in the MASTER dtype --
`storage_dtype` for weights (bfloat16 in the 32*_mixer configs, so a restart there resumes from bf16-rounded
weights) and the optimizer's master dtype for slots -- next to the global step.  Two containers hold the same names,
dtypes and contents: the TensorBundle pair + `checkpoint` state file that tf.train.Saver writes (tf_checkpoint.py,
restated without TensorFlow; `fmt="tf"`), and a single .npz (`fmt="npz"`).

    save_checkpoint(params, path)            after any number of steps
    load_checkpoint(params, path)            on a fresh ModelParameter, BEFORE the first step

`exact=True` stores the fp32 slices instead of the storage dtype: resuming is then bit-identical to not stopping."""
import os
import re
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


def _collect(params: ModelParameter, exact: bool) -> typing.Dict[str, torch.Tensor]:
    """name -> tensor in the dtype it is stored in (reference names, master dtypes; fp32 slices when `exact`)."""
    if not params.variable_cache:
        raise RuntimeError("nothing to save: the model has not been built yet")
    out = {"global_step": torch.tensor(params.global_step, dtype=torch.int64),
           "manual_global_step": torch.tensor(params.manual_global_step, dtype=torch.int64)}
    wdt = torch.float32 if exact else params.variable_dtype.master_dtype
    for name, var in params.variable_cache.items():
        out[name] = var.data.detach().to(wdt)
    state = getattr(params, "optimizer_state", None)
    if state is not None:
        sdt = params.optimizer_dtype.master_dtype
        for name, slot in state.slots.items():
            out[name] = slot.detach().to(slot.dtype if exact else sdt)
    return out


def save_checkpoint(params: ModelParameter, path: str, exact: bool = False, fmt: typing.Optional[str] = None) -> str:
    """fmt "npz": one .npz at `path`.  fmt "tf": `path` is the model directory; writes the TensorBundle pair
    `model.ckpt-<global_step>.index` / `.data-00000-of-00001` and the `checkpoint` state file, as tf.train.Saver does
    for the reference (src/run/run.py:158-175) -- see tf_checkpoint.py.  Default: "tf" for a directory (or a path
    without extension that is not an .npz), else "npz".  Returns the checkpoint's path / prefix."""
    tensors = _collect(params, exact)
    if fmt is None:
        fmt = "npz" if path.endswith(".npz") else ("tf" if os.path.isdir(path) else "npz")
    if fmt == "tf":
        from . import tf_checkpoint
        tensors = _to_graph_names(params, tensors)
        os.makedirs(path, exist_ok=True)
        name = f"model.ckpt-{params.global_step}"
        tf_checkpoint.write_bundle(os.path.join(path, name), tensors)
        older = []
        state_file = os.path.join(path, "checkpoint")
        if os.path.exists(state_file):
            with open(state_file) as f:
                older = [ln.split(":", 1)[1].strip().strip('"') for ln in f if ln.startswith("all_model_checkpoint_paths:")]
        tf_checkpoint.write_checkpoint_state(path, name, [o for o in older if o != name])
        return os.path.join(path, name)
    if fmt != "npz":
        raise ValueError(f"unknown checkpoint format '{fmt}'")
    out = {}
    for name, t in tensors.items():
        tag, arr = _pack(t, t.dtype)
        out[name + tag] = arr
    np.savez(path, **out)
    return path


def _to_graph_names(params: ModelParameter, tensors: typing.Dict[str, torch.Tensor]) -> typing.Dict[str, torch.Tensor]:
    """Slots under the names they have in the reference's graph (`<stage><k>v/...`, optimizer/backend.py:
    graph_slot_name) -- the keys a TF checkpoint of the reference holds; variables keep their names."""
    from ..optimizer.backend import graph_slot_name
    opt = f"/{params.optimizer.replace(':', '_')}/"
    order = {name: (k, len(var.data.shape)) for k, (name, var) in enumerate(params.variable_cache.items())}
    out = {}
    for name, t in tensors.items():
        base, sep, slot = name.rpartition(opt)
        if sep and base in order:
            k, rank = order[base]
            try:
                name = graph_slot_name(params, base, slot, k, rank)
            except KeyError:
                continue  # a buffer of this implementation that no stage of the optimizer string owns (all zeros)
        out[name] = t
    return out


def _slot_defined(params: ModelParameter, name: str) -> bool:
    """Does any stage of the optimizer string create a slot of this name (for some rank)?"""
    from ..optimizer.backend import slot_stage
    slot = name.rsplit("/", 1)[-1]
    for rank in (0, 4):
        try:
            slot_stage(params.optimizer, rank, slot)
            return True
        except KeyError:
            pass
    return False


_GRAPH_PREFIX = re.compile(r"^[a-z_]+\d+v/(.+)$")


def _from_graph_names(entries: typing.Dict[str, torch.Tensor]) -> typing.Dict[str, torch.Tensor]:
    """Inverse of _to_graph_names on read: `<stage><k>v/<variable>/<optimizer string>/<slot>` -> the internal slot
    name, recognised by its remainder starting with a variable that is itself in the file."""
    out = {}
    for name, t in entries.items():
        m = _GRAPH_PREFIX.match(name)
        if m:
            rest = m.group(1)
            parts = rest.split("/")
            # <variable>/<optimizer string>/<slot>: the variable is everything but the last two components
            if len(parts) > 2 and "/".join(parts[:-2]) in entries:
                name = rest
        out[name] = t
    return out


def _read_entries(path: str) -> typing.Dict[str, torch.Tensor]:
    """An .npz, a TensorBundle prefix, or a model directory with a `checkpoint` state file."""
    from . import tf_checkpoint
    if os.path.isdir(path):
        latest = tf_checkpoint.latest_checkpoint(path)
        if latest is None:
            raise FileNotFoundError(f"{path} holds no `checkpoint` state file")
        path = latest
    if os.path.exists(path + ".index"):
        entries = _from_graph_names(tf_checkpoint.read_bundle(path))
        return {k: (v.float() if v.dtype == torch.bfloat16 else v) for k, v in entries.items()}
    with np.load(path if path.endswith(".npz") else path + ".npz") as f:
        return dict(_unpack(k, f[k]) for k in f.files)


def load_checkpoint(params: ModelParameter, path: str):
    """Arms `params` so that variables (at creation) and optimizer slots (when the optimizer state is created) take
    their values from the file; restores the step counters at once.  A variable or slot the model asks for and the
    file lacks raises KeyError at that point (a file saved before the first update holds no slots at all: those
    start from zero, as in the reference); entries nobody asked for are reported by `finish_restore`, which the
    training step calls once everything exists, and which releases the host copy of the file."""
    entries = _read_entries(path)
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
        if kind == "slot" and (not (set(entries) - set(params.variable_cache)) or not _slot_defined(params, name)):
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
