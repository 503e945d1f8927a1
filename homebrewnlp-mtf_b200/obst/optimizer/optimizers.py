"""Optimizer stages chained by name (src/optimizer/optimizers.py:131-151).

The reference's stage functions add graph nodes; one `sess.run` later executes the whole chain per variable.
Here each stage appends itself to `ctx.plan`, and `lower_plan` maps the finished plan onto kernels:
    [adaptive_clip:<c>] - sm3 - momentum:<beta>:1:<nesterov> - learning_rate
(the hot path's chain) becomes ONE fused HBM pass (obst_opt_norms / obst_opt_update / obst_opt_finalize); any other
chain of registered stages runs stage by stage (obst_opt_stage / obst_opt_sm3_stage), one launch each."""
import typing

from .context import OptimizerCtx

# name -> number of string arguments the reference's stage function takes (graft: the inner optimizer + its own)
STAGE_ARGS = {"adam": 0, "sm3": 0, "novograd": 0, "adaptive_clip": 1, "l2norm_clip": 1, "value_clip": 1,
              "gradient_centralisation": 0, "weight_centralisation": 0, "learning_rate": 0, "global_l2norm_clip": 1,
              "momentum": 3}


def _check_args(name: str, args: typing.Tuple[str, ...]):
    if name == "graft":
        if not args or args[0] not in STAGE_ARGS:
            raise ValueError(f"graft needs a registered inner optimizer, got {args!r}")
        _check_args(args[0], tuple(args[1:]))
        return
    if len(args) != STAGE_ARGS[name]:
        raise TypeError(f"optimizer stage '{name}' takes {STAGE_ARGS[name]} argument(s), got {len(args)}: {args!r}")
    for a in args:
        float(a)


def _stage(name: str):
    def fn(ctx: OptimizerCtx, *args: str):
        _check_args(name, tuple(args))
        ctx.plan.append((name, tuple(args)))
        return ctx.grad

    fn.__name__ = name
    return fn


OPTIMIZERS = {name: _stage(name) for name in
              ("adam", "sm3", "novograd", "adaptive_clip", "l2norm_clip", "value_clip", "gradient_centralisation",
               "weight_centralisation", "learning_rate", "global_l2norm_clip", "momentum", "graft")}
sm3, momentum = OPTIMIZERS["sm3"], OPTIMIZERS["momentum"]
adaptive_gradient_clipping, multiply_learning_rate = OPTIMIZERS["adaptive_clip"], OPTIMIZERS["learning_rate"]


class FusedPlan(typing.NamedTuple):
    clip: float       # <= 0: no adaptive clipping
    beta: float
    nesterov: bool


class GenericPlan(typing.NamedTuple):
    stages: typing.Tuple[typing.Tuple[str, typing.Tuple[str, ...]], ...]


def lower_plan(plan: typing.List[typing.Tuple[str, tuple]]) -> typing.Union[FusedPlan, GenericPlan]:
    names = [p[0] for p in plan]
    clip = 0.0
    rest = list(plan)
    if names and names[0] == "adaptive_clip":
        clip = float(rest[0][1][0])
        rest = rest[1:]
    if [p[0] for p in rest] == ["sm3", "momentum", "learning_rate"]:
        margs = rest[1][1]
        if float(margs[1]) == 1.0:
            return FusedPlan(clip, float(margs[0]), bool(int(margs[2])))
    return GenericPlan(tuple((n, tuple(a)) for n, a in plan))
