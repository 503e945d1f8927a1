"""Optimizer stages chained by name (src/optimizer/optimizers.py:131-151).

The reference's stage functions add graph nodes; one `sess.run` later executes the whole chain per variable.
Here each stage appends itself to `ctx.plan`, and `lower_plan` maps the finished plan onto the fused multi-tensor
kernels (obst_opt_norms / obst_opt_update / obst_opt_finalize): the plan
    [adaptive_clip:<c>] - sm3 - momentum:<beta>:1:<nesterov> - learning_rate
is what those kernels realise in one HBM pass.  Other registered stages raise until they get a kernel
(SURVEY.md 8(f) item 1)."""
import typing

from .context import OptimizerCtx


def _stage(name: str):
    def fn(ctx: OptimizerCtx, *args: str):
        ctx.plan.append((name, tuple(args)))
        return ctx.grad

    fn.__name__ = name
    return fn


def _unsupported(name: str):
    def fn(ctx: OptimizerCtx, *args: str):
        raise NotImplementedError(f"optimizer stage '{name}' is registered by the reference but has no fused B200 "
                                  f"kernel yet (SURVEY.md 8(f))")

    fn.__name__ = name
    return fn


sm3 = _stage("sm3")
adaptive_gradient_clipping = _stage("adaptive_clip")
multiply_learning_rate = _stage("learning_rate")
momentum = _stage("momentum")

OPTIMIZERS = {"adam": _unsupported("adam"),
              "sm3": sm3,
              "novograd": _unsupported("novograd"),
              "adaptive_clip": adaptive_gradient_clipping,
              "l2norm_clip": _unsupported("l2norm_clip"),
              "value_clip": _unsupported("value_clip"),
              "gradient_centralisation": _unsupported("gradient_centralisation"),
              "weight_centralisation": _unsupported("weight_centralisation"),
              "learning_rate": multiply_learning_rate,
              "global_l2norm_clip": _unsupported("global_l2norm_clip"),
              "momentum": momentum,
              "graft": _unsupported("graft")}


class FusedPlan(typing.NamedTuple):
    clip: float       # <= 0: no adaptive clipping
    beta: float
    nesterov: bool


def lower_plan(plan: typing.List[typing.Tuple[str, tuple]]) -> FusedPlan:
    names = [p[0] for p in plan]
    clip = 0.0
    rest = list(plan)
    if names and names[0] == "adaptive_clip":
        clip = float(rest[0][1][0])
        rest = rest[1:]
    if [p[0] for p in rest] != ["sm3", "momentum", "learning_rate"]:
        raise NotImplementedError(f"optimizer chain {'-'.join(names)!r}: the fused kernel implements "
                                  f"'[adaptive_clip:c-]sm3-momentum:b:1:n-learning_rate'")
    margs = rest[1][1]
    if len(margs) != 3 or float(margs[1]) != 1.0:
        raise NotImplementedError(f"momentum:{':'.join(margs)}: gradient_multiplier must be 1")
    return FusedPlan(clip, float(margs[0]), bool(int(margs[2])))
