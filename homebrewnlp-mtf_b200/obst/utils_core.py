"""Deterministic scope naming, mirroring src/utils_core.py:16-19,57-67 of the reference.

The reference names every op/variable `prefix + counter` and nests TF variable scopes; variable identity
(src/utils_mtf.py:271-280) and the weight-decay predicate (src/optimizer/__init__.py:49-57) both hang off those
names, so the same scheme is kept here without TensorFlow: a module-level counter dict and a scope stack.
"""
import typing

NAME_INDICES: typing.Dict[str, int] = {}
_SCOPE: typing.List[str] = []


def random_name(prefix: str = "") -> str:
    """`prefix` + a per-prefix counter (the reference's "random" names are deterministic counters)."""
    if prefix not in NAME_INDICES:
        NAME_INDICES[prefix] = -1
    NAME_INDICES[prefix] += 1
    return f"{prefix}{NAME_INDICES[prefix]}"


def scoped(name: str, fn: typing.Callable, *args, **kwargs):
    name = random_name(name)
    _SCOPE.append(f"{name}v")
    try:
        return fn(*args, **kwargs)
    finally:
        _SCOPE.pop()


def current_scope() -> str:
    return "/".join(_SCOPE)


def scope_list() -> typing.List[str]:
    return list(_SCOPE)


def snapshot() -> typing.Dict[str, int]:
    """Counter state, so a reversible sub-block can be re-run under the names of its first execution."""
    return dict(NAME_INDICES)


def restore(state: typing.Dict[str, int]):
    NAME_INDICES.clear()
    NAME_INDICES.update(state)


def default(value: typing.Any, default_value: typing.Any) -> typing.Any:
    return default_value if value is None else value
