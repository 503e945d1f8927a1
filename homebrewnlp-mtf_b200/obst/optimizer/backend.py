"""Optimizer-state naming (src/optimizer/backend.py:23-25): `<var.name>/<optimizer ':'->'_'>/<slot>`, zeros,
non-trainable, optimizer_dtype."""
from ..dataclass import ModelParameter


def slot_name(params: ModelParameter, base_name: str, name: str) -> str:
    return f"{base_name}/{params.optimizer.replace(':', '_')}/{name}"


def slot_stage(optimizer: str, rank: int, slot: str) -> str:
    """The stage of the optimizer string that creates `slot` for a variable of that rank: sm3 owns dim<i> (and, for a
    rank-0 variable, adam's two moments: src/optimizer/optimizers.py:61-62), novograd / adam own exp_avg_p1 / p2,
    momentum owns momentum; `graft:<inner>` creates its inner optimizer's slots inside its own scope."""
    for stage in optimizer.split('-'):
        name, *args = stage.split(':')
        inner = args[0] if name == "graft" and args else name
        creates = {"sm3": ["exp_avg_p1", "exp_avg_p2"] if rank == 0 else [f"dim{i}" for i in range(rank)],
                   "adam": ["exp_avg_p1", "exp_avg_p2"], "novograd": ["exp_avg_p1", "exp_avg_p2"],
                   "momentum": ["momentum"]}.get(inner, [])
        if slot in creates:
            return name
    raise KeyError(f"no stage of '{optimizer}' creates the slot '{slot}'")


def graph_slot_name(params: ModelParameter, base_name: str, name: str, variable_index: int, rank: int) -> str:
    """The slot's name in the REFERENCE'S GRAPH, which is what tf.train.Saver writes: update() runs every stage inside
    `scoped(<stage>, ...)` (src/optimizer/__init__.py:42-44), whose name counter advances once per variable, so the
    slots of the k-th trainable variable live under `<stage><k>v/` (observed by running the reference:
    tests/golden/reference_optimizer.npz).  `slot_name` without that prefix is what this package uses internally."""
    stage = slot_stage(params.optimizer, rank, name)
    return f"{stage}{variable_index}v/{slot_name(params, base_name, name)}"
