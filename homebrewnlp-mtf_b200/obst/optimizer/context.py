"""Per-step bag handed to the optimizer stages (src/optimizer/context.py of the reference, reduced to what the
hot path reads)."""
import typing


class OptimizerCtx:
    def __init__(self, params, update_ops: list, learning_rate: float, step_count: int, assign: bool):
        self.params = params
        self.update_ops = update_ops
        self.learning_rate = learning_rate
        self.step_count = step_count
        self.assign = assign
        self.beta1 = params.opt_beta1
        self.beta2 = params.opt_beta2
        # filled while the optimizer string is walked: the stages, in order, that the fused kernel must realise
        self.plan: typing.List[typing.Tuple[str, tuple]] = []
        self.var = None
        self.grad = None
        self.value = None
