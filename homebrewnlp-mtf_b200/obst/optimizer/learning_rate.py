"""Learning-rate schedule modules (src/optimizer/learning_rate.py:27-72).  The reference evaluates them as TF
scalar ops on the float32 global step, read BEFORE it is incremented -- so with linear_warmup the very first
step runs with lr == 0.  Here they are host-side float32 arithmetic."""
import typing

import numpy as np

from ..dataclass import LearningRateConfig, ModelParameter

f32 = np.float32


class LearningRateCtx:
    def __init__(self, params: ModelParameter, loss_list: typing.Optional[list] = None,
                 update_ops: typing.Optional[list] = None):
        self.params = params
        self.learning_rate = f32(params.learning_rate)
        self.global_steps_float = f32(params.global_step)
        self.loss_list = loss_list
        self.global_steps_mtf = int(params.global_step)
        self.update_ops = update_ops
        self.config: typing.Optional[LearningRateConfig] = None


def linear_warmup(ctx: LearningRateCtx):
    warmup_steps_float = f32(ctx.config.final_step)
    is_warmup = f32(ctx.global_steps_float < warmup_steps_float)
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio = ctx.global_steps_float / warmup_steps_float
    warmup_factor = ratio * is_warmup + f32(1) * (f32(1) - is_warmup)  # weighted_add(ratio, 1, is_warmup)
    ctx.learning_rate = f32(ctx.learning_rate * warmup_factor)


def exponential_decay(ctx: LearningRateCtx):
    base = f32(ctx.config.factor)
    exp = np.maximum(ctx.global_steps_float - f32(ctx.config.start_step), f32(0))
    ctx.learning_rate = f32(ctx.learning_rate * np.power(base, exp))


def linear_decay(ctx: LearningRateCtx):
    start_step = f32(ctx.config.start_step)
    final_step = f32(ctx.config.final_step)
    current_step = ctx.global_steps_float - start_step
    final_step = final_step - start_step
    decay = f32(1) - current_step / final_step
    decay = np.maximum(np.minimum(decay, f32(1)), f32(0))
    ctx.learning_rate = f32(ctx.learning_rate * decay)


def lower_bound(ctx: LearningRateCtx):
    ctx.learning_rate = f32(np.maximum(ctx.learning_rate, f32(ctx.config.factor)))


def upper_bound(ctx: LearningRateCtx):
    ctx.learning_rate = f32(np.minimum(ctx.learning_rate, f32(ctx.config.factor)))


MODULES = {"linear_warmup": linear_warmup,
           "exponential_decay": exponential_decay,
           "linear_decay": linear_decay,
           "lower_bound": lower_bound,
           "upper_bound": upper_bound}


def get_learning_rate(params: ModelParameter, loss_list: typing.Optional[list] = None,
                      update_ops: typing.Optional[list] = None) -> LearningRateCtx:
    ctx = LearningRateCtx(params, loss_list, update_ops)
    for name, keys in params.learning_rate_config.items():
        ctx.config = keys
        MODULES[name](ctx)
    return ctx
