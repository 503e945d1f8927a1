"""get_optimizer: reverse-mode sweep + per-variable update (src/optimizer/__init__.py of the reference).

The reference differentiates the graph op by op (`:128-174`) and then emits ~25 small ops per variable
(`update`, `:31-66`).  Here the sweep replays the tape recorded by build(); the update of ALL variables is three
kernel launches (norms, update, finalize) over a device-side table of tensors.  Gradients are summed over
micro-batches and data-parallel ranks in one flat fp32 buffer before the update, so every rank applies the
identical step (adaptive-clip norms are taken after the allreduce)."""
import ctypes
import typing

import torch

from .. import native
from ..dataclass import ModelParameter
from ..model import backward as _backward
from ..model.backend import Variable
from ..utils_mtf import feature_dims_used
from .backend import slot_name
from .context import OptimizerCtx
from .learning_rate import get_learning_rate
from .optimizers import OPTIMIZERS, FusedPlan, GenericPlan, lower_plan

ITEM_ELEMS = 32768  # elements per work item of the fused kernels


def is_large_tensor(params: ModelParameter, var: Variable) -> bool:
    """The weight-decay predicate of src/optimizer/__init__.py:49-57, evaluated on the variable's scope name."""
    name = var.name
    features_used = feature_dims_used(params, var.dims)
    large_tensor = features_used and len(var.dims) > len(params.feature_dims)
    large_tensor |= not features_used and len(var.dims) >= 2
    large_tensor &= var.size > 1
    large_tensor &= "norm" not in name
    large_tensor &= "rezero" not in name
    large_tensor &= "embed" not in name
    large_tensor &= "input" not in name or "lang_in" in name or "vid_in" in name
    large_tensor &= "output" not in name or "lang_out" in name or "vid_out" in name
    return bool(large_tensor)


def _depth_of(var: Variable) -> int:
    """Depth index from the scope name `.../body0v/<depth>_<cfg>v/...`; -1 for input/output variables."""
    parts = var.name.split('/')
    if "body0v" in parts:
        blk = parts[parts.index("body0v") + 1]
        return int(blk.split('_')[0])
    return -1


class FusedOptimizer:
    """State + launch plumbing of the fused update.  Owns the flat gradient buffer (so DP reduction is one or a few
    contiguous allreduces), the momentum / SM3 slots and the device-side tensor table."""

    def __init__(self, params: ModelParameter, plan: FusedPlan):
        self.params = params
        self.plan = plan
        dev = torch.device(params.device)
        self.state_dtype = params.optimizer_dtype.slice_dtype
        self.act_dtype = params.variable_dtype.activation_dtype
        variables = [v for v in params.variable_cache.values() if v.trainable]
        if not variables:
            raise RuntimeError("no trainable variables: run build() once before get_optimizer()")
        for v in variables:
            if not 1 <= len(v.dims) <= 4:
                raise NotImplementedError(f"sm3 on rank-{len(v.dims)} variable {v.name}: ranks 1..4 are fused "
                                          f"(rank 0 falls back to adam in the reference, unused by the mixer)")
        # flat layout: per-depth variables first (their gradients are final first during the last backward pass),
        # shared mixing maps / input / output variables last
        shared = {id(v) for v in variables if _depth_of(v) == 0 and "embed" in v.name}
        order = sorted(range(len(variables)),
                       key=lambda i: (1 if (_depth_of(variables[i]) < 0 or id(variables[i]) in shared) else 0,
                                      -_depth_of(variables[i]) if _depth_of(variables[i]) >= 0 else 0, i))
        self.variables = [variables[i] for i in order]
        sizes = [v.size for v in self.variables]
        pad = [(-s) % 64 for s in sizes]  # keep every tensor 256-byte aligned inside the flat buffers
        total = sum(s + p for s, p in zip(sizes, pad))
        self.flat_grad = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_momentum = torch.zeros(total, dtype=self.state_dtype, device=dev)
        n_acc = sum(sum(d.size for d in v.dims) + 64 * len(v.dims) for v in self.variables)
        self.flat_acc = torch.zeros(n_acc, dtype=self.state_dtype, device=dev)
        self.flat_acc_new = torch.zeros(n_acc, dtype=torch.float32, device=dev)
        self.norms = torch.zeros(2 * len(self.variables), dtype=torch.float64, device=dev)
        self.slots: typing.Dict[str, torch.Tensor] = {}
        self.offsets: typing.List[typing.Tuple[int, int]] = []
        self.depth_ranges: typing.Dict[int, typing.List[int]] = {}

        table = (native.OptTensor * len(self.variables))()
        items = []  # (class, tensor, row0, rows)
        off = 0
        aoff = 0
        max_cols, vec4 = 1, True
        for ti, (v, sz, pd) in enumerate(zip(self.variables, sizes, pad)):
            old_grad = v.grad
            v.grad = self.flat_grad[off:off + sz].view(v.data.shape)
            if old_grad is not None:
                v.grad.copy_(old_grad)
            m = self.flat_momentum[off:off + sz].view(v.data.shape)
            self.slots[slot_name(params, v.name, "momentum")] = m
            self.offsets.append((off, sz))
            d = _depth_of(v)
            key = -1 if (d < 0 or id(v) in shared) else d
            rng = self.depth_ranges.setdefault(key, [off, off + sz + pd])
            rng[0] = min(rng[0], off)
            rng[1] = max(rng[1], off + sz + pd)
            t = table[ti]
            t.p = v.data.data_ptr()
            t.shadow = v.value.data_ptr() if v.value is not v.data else None
            t.g = v.grad.data_ptr()
            t.m = m.data_ptr()
            dims = [dd.size for dd in v.dims]
            for i, n in enumerate(dims):
                a = self.flat_acc[aoff:aoff + n]
                an = self.flat_acc_new[aoff:aoff + n]
                self.slots[slot_name(params, v.name, f"dim{i}")] = a
                t.acc[i] = a.data_ptr()
                t.acc_new[i] = an.data_ptr()
                t.dims[i] = n
                aoff += n + ((-n) % 64)
            t.numel = sz
            t.ndim = len(dims)
            t.decay = int(is_large_tensor(params, v))
            t.tril_s = v.tril_s
            cols = dims[-1]
            rows = sz // cols
            max_cols = max(max_cols, cols)
            vec4 &= cols % 4 == 0
            # launch class = the register footprint the update kernel needs for this last dim (columns per lane):
            # tensors of a class share one launch, so narrow tensors are not run with the widest instantiation
            per_lane = -(-cols // (128 if cols % 4 == 0 else 32))
            klass = 4 if per_lane <= 4 else (8 if per_lane <= 8 else (16 if per_lane <= 16 else 0))
            # the update kernel splits wide rows into 512-column chunks (one block each): size items per chunk
            rpi = min(2048, max(1, ITEM_ELEMS // min(cols, 512)))
            for r0 in range(0, rows, rpi):
                items.append((klass if cols % 4 == 0 else -klass - 1, ti, r0, min(rpi, rows - r0), cols))
            off += sz + pd
        items.sort(key=lambda it: (it[0], it[1], it[2]))
        items_t = [it[1] for it in items]
        items_r0 = [it[2] for it in items]
        items_rows = [it[3] for it in items]
        self.groups: typing.List[typing.Dict[str, int]] = []
        for pos, it in enumerate(items):
            if not self.groups or self.groups[-1]["key"] != it[0]:
                self.groups.append({"key": it[0], "item0": pos, "items": 0, "max_cols": 1, "vec4": int(it[0] >= 0)})
            g = self.groups[-1]
            g["items"] += 1
            g["max_cols"] = max(g["max_cols"], it[4])
        self.max_cols, self.vec4 = max_cols, vec4
        raw = bytes(table)
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.n_items = len(items_t)
        self.item_tensor = torch.tensor(items_t, dtype=torch.int32, device=dev)
        self.item_row0 = torch.tensor(items_r0, dtype=torch.int32, device=dev)
        self.item_rows = torch.tensor(items_rows, dtype=torch.int32, device=dev)
        self.decay_names = [v.name for v in self.variables if is_large_tensor(params, v)]
        # sum cast_AT(p)^2 per tensor, accumulated by the update itself for the NEXT step's adaptive-clip norm
        self.wnext = torch.zeros(len(self.variables), dtype=torch.float64, device=dev)
        self.wnorm_valid = False

    @property
    def n_params(self) -> int:
        return sum(sz for _, sz in self.offsets)

    def zero_grad(self):
        self.flat_grad.zero_()

    def invalidate_weight_norms(self):
        """Call after writing Variable.data from outside the optimizer (weight import): the carried norms are stale."""
        self.wnorm_valid = False
        self.wnext.zero_()

    # ------------------------------------------------------------------------------------------------------
    # generic chains: one launch per stage (obst_opt_stage), slots in flat buffers laid out like flat_grad
    # ------------------------------------------------------------------------------------------------------
    def _flat_state(self, slot: str) -> torch.Tensor:
        """Flat state buffer of one slot name for ALL variables (zeros on first use, src/optimizer/backend.py:23-25);
        per-variable views are published in self.slots under the reference's names."""
        if slot == "momentum":
            return self.flat_momentum
        bufs = self.__dict__.setdefault("_state_bufs", {})
        if slot not in bufs:
            buf = torch.zeros_like(self.flat_momentum)
            bufs[slot] = buf
            entries = getattr(self.params, "restore_entries", None) or {}  # resume (obst.run.checkpoint)
            for v, (off, sz) in zip(self.variables, self.offsets):
                name = slot_name(self.params, v.name, slot)
                view = buf[off:off + sz].view(v.data.shape)
                self.slots[name] = view
                if name in entries:
                    view.copy_(entries[name].to(view.device).reshape(view.shape).to(view.dtype))
        return bufs[slot]

    def _scratch(self, name: str, numel: int, dtype) -> torch.Tensor:
        bufs = self.__dict__.setdefault("_scratch_bufs", {})
        if name not in bufs:
            bufs[name] = torch.zeros(numel, dtype=dtype, device=self.flat_grad.device)
        return bufs[name]

    def _stage(self, stage: int, work: torch.Tensor, f=(), flags: int = 0, work2=None, state_a=None, state_b=None,
               scalars=None, sums=None, sums2=None):
        a = native.OptStageArgs()
        a.stage, a.flags = stage, flags
        a.dtype_act, a.dtype_state = native.dtype_code(self.act_dtype), native.dtype_code(self.state_dtype)
        for i, v in enumerate(f):
            a.f[i] = float(v)
        a.grad_base, a.work = self.flat_grad.data_ptr(), work.data_ptr()
        a.work2 = None if work2 is None else work2.data_ptr()
        a.state_a = None if state_a is None else state_a.data_ptr()
        a.state_b = None if state_b is None else state_b.data_ptr()
        if scalars is not None:
            a.scalars_in, a.scalars_out = scalars[0].data_ptr(), scalars[1].data_ptr()
        a.sums = None if sums is None else sums.data_ptr()
        a.sums2 = None if sums2 is None else sums2.data_ptr()
        native.check(native.load().obst_opt_stage(self.table.data_ptr(), len(self.variables),
                                                  self.item_tensor.data_ptr(), self.item_row0.data_ptr(),
                                                  self.item_rows.data_ptr(), self.n_items, ctypes.byref(a),
                                                  native.stream_ptr()), "obst_opt_stage")

    def _run_stage(self, name: str, args: typing.Tuple[str, ...], work: torch.Tensor, lr: float, step_count: int):
        """One registered stage (src/optimizer/optimizers.py:32-151) on the working gradient buffer `work`."""
        import numpy as np
        N = native
        nt = len(self.variables)
        f32 = np.float32
        b1, b2 = f32(self.params.opt_beta1), f32(self.params.opt_beta2)
        sums = self._scratch("sums", 4 * nt, torch.float64)

        def reduce_into(dst, src=work):
            self._stage(N.STAGE_REDUCE, src, flags=2, sums=dst)

        def debias(beta):  # debias_momentum, optimizers.py:17-18, in fp32 like the graph
            return f32(1) / (f32(1) - np.power(beta, f32(step_count)))

        if name == "adaptive_clip":
            reduce_into(sums)
            self._stage(N.STAGE_ADAPTIVE_CLIP, work, f=[args[0]], sums=sums)
        elif name == "l2norm_clip":
            reduce_into(sums)
            self._stage(N.STAGE_L2NORM_CLIP, work, f=[args[0]], sums=sums)
        elif name == "global_l2norm_clip":  # norm of the RAW gradients of all variables, taken once per step
            self._stage(N.STAGE_GLOBAL_L2NORM_CLIP, work, f=[args[0]], sums=self._scratch("sums_raw", 4 * nt,
                                                                                          torch.float64))
        elif name == "value_clip":
            self._stage(N.STAGE_VALUE_CLIP, work, f=[args[0]])
        elif name == "gradient_centralisation":
            reduce_into(sums)
            self._stage(N.STAGE_GRAD_CENTRALISE, work, sums=sums)
        elif name == "weight_centralisation":
            reduce_into(sums)
            self._stage(N.STAGE_WEIGHT_CENTRALISE, work, sums=sums)
        elif name == "learning_rate":
            self._stage(N.STAGE_LEARNING_RATE, work, f=[lr])
        elif name == "momentum":
            self._stage(N.STAGE_MOMENTUM, work, f=[args[0], args[1]], flags=int(bool(int(args[2]))),
                        state_a=self._flat_state("momentum"))
        elif name == "adam":
            self._stage(N.STAGE_ADAM, work, f=[b1, b2, debias(b1), debias(b2)],
                        state_a=self._flat_state("exp_avg_p1"), state_b=self._flat_state("exp_avg_p2"))
        elif name == "novograd":
            reduce_into(sums)
            if "_novograd_p2" not in self.__dict__:
                pair = [torch.zeros(nt, dtype=self.state_dtype, device=work.device) for _ in range(2)]
                entries = getattr(self.params, "restore_entries", None) or {}
                for i, v in enumerate(self.variables):
                    name = slot_name(self.params, v.name, "exp_avg_p2")
                    if name in entries:
                        pair[0][i] = entries[name].to(work.device).reshape(()).to(self.state_dtype)
                self._novograd_p2 = pair
            pair = self._novograd_p2
            self._stage(N.STAGE_NOVOGRAD, work, f=[b1, b2, 0.0, debias(b2)], state_a=self._flat_state("exp_avg_p1"),
                        scalars=pair, sums=sums)
            pair.reverse()  # [0] is always the current exp_avg_p2
            for i, v in enumerate(self.variables):
                self.slots[slot_name(self.params, v.name, "exp_avg_p2")] = pair[0][i]
        elif name == "sm3":
            table = self.table if work is self.flat_grad else self._shifted_table(work)
            lib, st, sdt = native.load(), native.stream_ptr(), native.dtype_code(self.state_dtype)
            for g in self.groups:
                o = 4 * g["item0"]
                native.check(lib.obst_opt_sm3_stage(table.data_ptr(), nt, self.item_tensor.data_ptr() + o,
                                                    self.item_row0.data_ptr() + o, self.item_rows.data_ptr() + o,
                                                    g["items"], g["max_cols"], g["vec4"], sdt, st),
                             "obst_opt_sm3_stage")
            native.check(lib.obst_opt_finalize(table.data_ptr(), nt, sdt, st), "obst_opt_finalize")
        elif name == "graft":  # direction of g, magnitude of the inner optimizer's step (optimizers.py:147-151)
            sums_g = self._scratch("sums_graft", 4 * nt, torch.float64)
            tmp = self._scratch("graft_work", self.flat_grad.numel(), torch.float32)
            reduce_into(sums_g)
            self._stage(N.STAGE_COPY, work, work2=tmp)
            self._run_stage(args[0], tuple(args[1:]), tmp, lr, step_count)
            sums_inner = self._scratch("sums_inner", 4 * nt, torch.float64)
            reduce_into(sums_inner, tmp)
            self._stage(N.STAGE_GRAFT_COMBINE, work, sums=sums_g, sums2=sums_inner)
        else:
            raise KeyError(name)

    def _shifted_table(self, work: torch.Tensor) -> torch.Tensor:
        """The tensor table with every g pointer moved into `work` (same layout as flat_grad): for the SM3 stage
        when it runs on a copy of the gradient (graft:sm3)."""
        cache = self.__dict__.setdefault("_shifted_tables", {})
        key = work.data_ptr()
        if key not in cache:
            n = len(self.variables)
            raw = bytearray(self.table.cpu().numpy().tobytes())
            table = (native.OptTensor * n).from_buffer(raw)
            delta = work.data_ptr() - self.flat_grad.data_ptr()
            for t in table:
                t.g = t.g + delta
            cache[key] = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).to(work.device)
        return cache[key]

    def _step_generic(self, lr: float, step_count: int, refresh_transposed: bool):
        N = native
        plan: GenericPlan = self.plan
        work = self.flat_grad
        if self.act_dtype == torch.bfloat16:  # the graph hands over activation-dtype gradients
            self._stage(N.STAGE_ROUND_BF16, work)

        def uses(stage_name, stages):
            return any(n == stage_name or (n == "graft" and a and a[0] == stage_name) for n, a in stages)

        if uses("global_l2norm_clip", plan.stages):
            self._stage(N.STAGE_REDUCE, work, flags=2, sums=self._scratch("sums_raw", 4 * len(self.variables),
                                                                          torch.float64))
        for name, args in plan.stages:
            self._run_stage(name, args, work, lr, step_count)
        self._stage(N.STAGE_APPLY, work, f=[float(lr) * float(self.params.weight_decay)])
        self.wnorm_valid = False
        if refresh_transposed:
            for v in self.variables:
                v.refresh_t()

    def step(self, lr: float, refresh_transposed: bool = True, step_count: int = 1):
        if isinstance(self.plan, GenericPlan):
            return self._step_generic(lr, step_count, refresh_transposed)
        lib = native.load()
        st = native.stream_ptr()
        p, plan = self.params, self.plan
        nt = len(self.variables)
        grad_bf16 = int(self.act_dtype == torch.bfloat16)
        act, sdt = native.dtype_code(self.act_dtype), native.dtype_code(self.state_dtype)
        tp, it, ir0, irs = (self.table.data_ptr(), self.item_tensor.data_ptr(), self.item_row0.data_ptr(),
                            self.item_rows.data_ptr())
        carry = 1 if self.wnorm_valid else 0
        if plan.clip > 0:  # one pass over every gradient (and, on the first step only, every weight)
            native.check(lib.obst_opt_norms(tp, nt, it, ir0, irs, self.n_items, act, grad_bf16,
                                            self.norms.data_ptr(), carry, self.wnext.data_ptr(), 0, nt, st),
                         "obst_opt_norms")
        for g in self.groups:
            o = 4 * g["item0"]
            native.check(lib.obst_opt_update(tp, nt, it + o, ir0 + o, irs + o, g["items"], g["max_cols"], g["vec4"],
                                             act, sdt, grad_bf16, self.norms.data_ptr(), self.wnext.data_ptr(),
                                             plan.clip, plan.beta, int(plan.nesterov), float(lr),
                                             float(p.weight_decay), st), "obst_opt_update")
        native.check(lib.obst_opt_finalize(tp, nt, sdt, st), "obst_opt_finalize")
        if refresh_transposed:  # the update rewrote the shadows in place: transposed copies follow
            for v in self.variables:
                v.refresh_t()
        self.wnorm_valid = plan.clip > 0


def update(ctx: OptimizerCtx):
    """src/optimizer/__init__.py:31-66 for ALL variables at once: walk the optimizer string, lower the plan, run."""
    params = ctx.params
    ctx.plan = []
    for opt in params.optimizer.split('-'):
        opt, *args = opt.split(':')
        ctx.grad = OPTIMIZERS[opt](ctx, *args)
    plan = lower_plan(ctx.plan)
    opt_state: FusedOptimizer = params.optimizer_state
    if opt_state.plan != plan:
        raise RuntimeError("optimizer string changed after the optimizer state was created")
    opt_state.step(ctx.learning_rate, step_count=ctx.step_count)
    ctx.update_ops.extend(v.name for v in opt_state.variables)


def _ensure_state(params: ModelParameter) -> FusedOptimizer:
    state = getattr(params, "optimizer_state", None)
    if state is None:
        ctx = OptimizerCtx(params, [], 0.0, 0, True)
        for opt in params.optimizer.split('-'):
            opt, *args = opt.split(':')
            OPTIMIZERS[opt](ctx, *args)
        state = FusedOptimizer(params, lower_plan(ctx.plan))
        params.optimizer_state = state
        if getattr(params, "restore_entries", None) is not None:  # resume: slots come from the checkpoint
            from ..run.checkpoint import restore_slots
            restore_slots(params, state)
    return state


def get_optimizer(loss_list: typing.List[typing.Any], params: ModelParameter, manual_step: int, fn: str
                  ) -> typing.Tuple[typing.List[str], float]:
    """Differentiate the recorded step and, when fn == "update", apply the parameter update.
    Same contract as the reference (src/optimizer/__init__.py:69-186): returns (update_ops, learning_rate);
    any fn other than "update" only accumulates gradients (the reference's "add"/"accumulate" arm, used for all
    but the last micro-batch)."""
    state = _ensure_state(params)
    update_ops: typing.List[str] = []
    learning_rate_ctx = get_learning_rate(params, loss_list, update_ops)
    learning_rate = float(learning_rate_ctx.learning_rate)
    reducer = getattr(params, "grad_reducer", None)
    if reducer is not None:
        reducer.begin_backward(last=(fn == "update"))
    _backward(params)
    if fn != "update":
        return update_ops, learning_rate
    if reducer is not None:
        reducer.finish()
    ctx = OptimizerCtx(params, update_ops, learning_rate, int(params.global_step) + 1, True)
    update(ctx)
    return ctx.update_ops, learning_rate
