"""Optimizer-state naming (src/optimizer/backend.py:23-25): `<var.name>/<optimizer ':'->'_'>/<slot>`, zeros,
non-trainable, optimizer_dtype."""
from ..dataclass import ModelParameter


def slot_name(params: ModelParameter, base_name: str, name: str) -> str:
    return f"{base_name}/{params.optimizer.replace(':', '_')}/{name}"
