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

ITEM_ELEMS = 131072  # elements of one 512-column chunk per work item of the fused kernels (32 rows per warp)


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


def _drain_list(items, cols_of, slots: int = 296):
    """Work list of the update kernel.  Blocks are dispatched in list order onto `slots` resident blocks (148 SMs x 2);
    with items of one size the last wave leaves most of them idle (2612 equal items = 8.8 waves: 2 % of the kernel).
    The first part of the list keeps the full-size items (long rows runs stream best); the last stretch is cut
    finer and finer, so that every slot runs dry at about the same time.
    OBST_OPT_DRAIN = "<tail fraction>:<divisor 1>:<divisor 2>" overrides the grading (0 = off)."""
    import os
    frac, div1, div2 = (float(x) for x in os.environ.get("OBST_OPT_DRAIN", "0.05:3:6").split(":"))
    if frac <= 0:
        return list(items)
    work = [rows * cols_of[ti] for ti, _, rows in items]
    total = float(sum(work)) or 1.0
    out, done = [], 0.0
    for (ti, r0, rows), w in zip(items, work):
        pos = done / total
        done += w
        div = 1 if pos < 1 - frac else (div1 if pos < 1 - frac / 3 else div2)
        step = max(8, -(-rows // int(div)))  # a warp per row: never fewer rows than warps
        for a in range(0, rows, step):
            out.append((ti, r0 + a, min(step, rows - a)))
    return out


def _depth_of(var: Variable) -> int:
    """Depth index from the scope name `.../body0v/<depth>_<cfg>v/...`; -1 for input/output variables."""
    import re
    parts = var.name.split('/')
    if "body0v" in parts:
        # the block scope follows body0v directly, or one level deeper under the 'checkpoint' strategy
        # (body0v/recompute_grad<n>v/<depth>_<cfg>v); body-level variables that belong to no block (the initial
        # position embedding) go with the input / output bucket
        for blk in parts[parts.index("body0v") + 1:][:2]:
            m = re.fullmatch(r"(\d+)_\d+v", blk)
            if m:
                return int(m.group(1))
    return -1


class FusedOptimizer:
    """State + launch plumbing of the fused update.  Owns the flat gradient buffer (so DP reduction is one or a few
    contiguous allreduces), the momentum / SM3 slots and the device-side tensor table."""

    def __init__(self, params: ModelParameter, plan: FusedPlan, variables: typing.Optional[list] = None):
        self.params = params
        self.plan = plan
        self.scalar_opt: typing.Optional["FusedOptimizer"] = None
        dev = torch.device(params.device)
        self.state_dtype = params.optimizer_dtype.slice_dtype
        self.act_dtype = params.variable_dtype.activation_dtype
        # The graph hands the optimizer gradients in the activation dtype (src/optimizer/__init__.py:180).  For the
        # fused chain with bf16 activations they are therefore kept as a second, bf16 flat buffer -- written once per
        # step from the fp32 accumulation buffer (obst_grad_finalize), all-reduced in that dtype, read by the update.
        # Generic chains rewrite the gradient stage by stage and stay on the fp32 buffer.
        self.grad_lp = self.act_dtype == torch.bfloat16 and isinstance(plan, FusedPlan)
        if variables is None:
            variables = [v for v in params.variable_cache.values() if v.trainable]
            if not variables:
                raise RuntimeError("no trainable variables: run build() once before get_optimizer()")
            scalars = [v for v in variables if len(v.dims) == 0]
            variables = [v for v in variables if len(v.dims) > 0]
            if scalars:
                # rank-0 variables (rezero gates): sm3 / novograd fall back to adam for them
                # (src/optimizer/optimizers.py:48-49,61-62) -- a second, stage-by-stage state holds them
                self.scalar_opt = FusedOptimizer(params, _scalar_plan(plan), variables=scalars)
        for v in variables:
            if len(v.dims) > 4:
                raise NotImplementedError(f"sm3 on rank-{len(v.dims)} variable {v.name}: ranks 0..4 are supported")
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
        self.flat_grad_lp = torch.zeros(total, dtype=torch.bfloat16, device=dev) if self.grad_lp else None
        self.flat_momentum = torch.zeros(total, dtype=self.state_dtype, device=dev)
        n_acc = sum(sum(d.size for d in v.dims) + 64 * len(v.dims) for v in self.variables)
        self.flat_acc = torch.zeros(n_acc, dtype=self.state_dtype, device=dev)
        self.flat_acc_new = torch.zeros(n_acc, dtype=torch.float32, device=dev)
        self.norms = torch.zeros(2 * len(self.variables), dtype=torch.float64, device=dev)
        self.slots: typing.Dict[str, torch.Tensor] = {}
        self.offsets: typing.List[typing.Tuple[int, int]] = []
        self.depth_ranges: typing.Dict[int, typing.List[int]] = {}
        self.item_ranges: typing.Dict[int, typing.List[int]] = {}  # depth key -> [first item, end item)
        self.finalized: typing.Set[int] = set()                     # depth keys handed over this step
        self.g2_keys: typing.Set[int] = set()                       # ... whose sum g^2 the hand-over already took

        table = (native.OptTensor * len(self.variables))()
        items = []  # (tensor, row0, rows), in flat-layout order
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
            if self.grad_lp:
                t.g = self.flat_grad_lp.data_ptr() + 2 * off
                t.g_acc = v.grad.data_ptr()
                t.g_bf16 = 1
            else:
                t.g = v.grad.data_ptr()
                t.g_acc = None
                t.g_bf16 = 0
            t.m = m.data_ptr()
            dims = [dd.size for dd in v.dims] or [1]  # a rank-0 variable is laid out as one element
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
            # the update kernels split wide rows into 512-column chunks (one block each): size items per chunk
            rpi = min(2048, max(1, ITEM_ELEMS // min(cols, 512)))
            irng = self.item_ranges.setdefault(key, [len(items), len(items)])
            if irng[1] != len(items):
                raise RuntimeError("flat layout is not grouped by depth")  # the sort above guarantees it
            for r0 in range(0, rows, rpi):
                items.append((ti, r0, min(rpi, rows - r0)))
            irng[1] = len(items)
            off += sz + pd
        self.max_cols, self.vec4 = max_cols, vec4
        raw = bytes(table)
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.n_items = len(items)
        self.item_tensor = torch.tensor([it[0] for it in items], dtype=torch.int32, device=dev)
        self.item_row0 = torch.tensor([it[1] for it in items], dtype=torch.int32, device=dev)
        self.item_rows = torch.tensor([it[2] for it in items], dtype=torch.int32, device=dev)
        # the update kernel's own work list: same tensors, items graded in size towards the end (see _drain_list)
        upd = _drain_list(items, [int(table[ti].dims[table[ti].ndim - 1]) for ti in range(len(self.variables))])
        self.n_update_items = len(upd)
        self.update_items = tuple(torch.tensor([it[k] for it in upd], dtype=torch.int32, device=dev) for k in range(3))
        self.decay_names = [v.name for v in self.variables if is_large_tensor(params, v)]
        # sum cast_AT(p)^2 per tensor, accumulated by the update itself for the NEXT step's adaptive-clip norm
        self.wnext = torch.zeros(len(self.variables), dtype=torch.float64, device=dev)
        self.wnorm_valid = False
        self._t_sig = None   # (shadow, transposed shadow) addresses the transpose table was built for
        self._t_table = None
        self._t_tiles = 0
        if self.scalar_opt is not None:
            self.slots.update(self.scalar_opt.slots)

    @property
    def n_params(self) -> int:
        return sum(sz for _, sz in self.offsets)

    def zero_grad(self):
        """Start of a step: clear the fp32 accumulation buffer, the norm sums and the hand-over bookkeeping."""
        self.flat_grad.zero_()
        self.norms.zero_()
        self.finalized.clear()
        self.g2_keys.clear()
        if self.scalar_opt is not None:
            self.scalar_opt.zero_grad()

    def extra_reduce_buffers(self) -> typing.List[torch.Tensor]:
        """Gradient buffers outside the depth buckets that data-parallel ranks must also sum (rank-0 variables)."""
        return [] if self.scalar_opt is None else [self.scalar_opt.flat_grad]

    def reduce_buffer(self) -> torch.Tensor:
        """What data-parallel ranks all-reduce: the gradients in the dtype the optimizer is handed."""
        return self.flat_grad_lp if self.grad_lp else self.flat_grad

    def hand_over(self, key: typing.Optional[int], with_norms: bool):
        """g(bf16) <- round(fp32 accumulation) for one depth bucket (`key`; None = every bucket not yet handed
        over): the cast of src/optimizer/__init__.py:180 at the point where the bucket's gradient is final.
        with_norms: also sum the squares of the rounded values (single rank: nothing is reduced afterwards)."""
        if not self.grad_lp:
            return
        keys = [k for k in self.item_ranges if k not in self.finalized] if key is None else [key]
        lib, st = native.load(), native.stream_ptr()
        # contiguous item ranges are merged into one launch
        spans = sorted(tuple(self.item_ranges[k]) for k in keys if k not in self.finalized)
        merged: typing.List[typing.List[int]] = []
        for a, b in spans:
            if merged and merged[-1][1] == a:
                merged[-1][1] = b
            elif b > a:
                merged.append([a, b])
        for a, b in merged:
            native.check(lib.obst_grad_finalize(self.table.data_ptr(), len(self.variables),
                                                self.item_tensor.data_ptr() + 4 * a, self.item_row0.data_ptr() + 4 * a,
                                                self.item_rows.data_ptr() + 4 * a, b - a, self.norms.data_ptr(),
                                                int(with_norms), st), "obst_grad_finalize")
        self.finalized.update(keys)
        if with_norms:
            self.g2_keys.update(keys)

    def refresh_transposed(self):
        """Rewrite every transposed weight shadow from the shadow the update kernel just wrote: one launch."""
        # weights with a group norm folded into them (basic._mid_norm_out_folded): one launch per shape
        folded = [v for v in self.variables if v.folded is not None]
        if folded:
            sig = tuple((v.value.data_ptr(), v.folded[5].data_ptr()) for v in folded)
            if sig != self.__dict__.get("_fold_sig"):
                groups: typing.Dict[tuple, list] = {}
                for v in folded:
                    groups.setdefault(tuple(v.folded[2:5]), []).append(v.fold_entry())
                self._fold_tables = [(native.fold_items(entries), len(entries), shape)
                                     for shape, entries in groups.items()]
                self._fold_sig = sig
            for table, n, (heads, kin, kout) in self._fold_tables:
                native.fold_norm_weight(table, n, heads, kin, kout)
        pairs = [(v, v.value, v.value_t) for v in self.variables if v.value_t is not None]
        if not pairs:
            return
        sig = tuple((a.data_ptr(), b.data_ptr()) for _, a, b in pairs)
        if sig != self._t_sig:
            items = (native.TransposeItem * len(pairs))()
            tiles = 0
            for it, (v, a, b) in zip(items, pairs):
                batch, rows, cols = v._t_view
                it.src, it.dst = a.data_ptr(), b.data_ptr()
                it.batch, it.rows, it.cols, it.tile0 = batch, rows, cols, tiles
                tiles += batch * (-(-rows // 64)) * (-(-cols // 64))
            self._t_table = torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8).to(self.table.device)
            self._t_sig, self._t_tiles, self._t_n = sig, tiles, len(pairs)
        native.check(native.load().obst_transpose_batched(native.dtype_code(self.act_dtype), self._t_table.data_ptr(),
                                                          self._t_n, self._t_tiles, native.stream_ptr()),
                     "obst_transpose_batched")

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
            from ..run.checkpoint import restore_slot  # resume: no-op unless load_checkpoint armed the parameters
            for v, (off, sz) in zip(self.variables, self.offsets):
                name = slot_name(self.params, v.name, slot)
                view = buf[off:off + sz].view(v.data.shape)
                self.slots[name] = view
                restore_slot(self.params, name, view)
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
            reducer = getattr(self.params, "grad_reducer", None)
            if reducer is not None:  # replicas agree on the sums (see GradReducer.agree)
                reducer.agree(dst)

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
                from ..run.checkpoint import restore_slot
                for i, v in enumerate(self.variables):
                    restore_slot(self.params, slot_name(self.params, v.name, "exp_avg_p2"), pair[0][i])
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
            native.check(lib.obst_opt_sm3_stage(table.data_ptr(), nt, self.item_tensor.data_ptr(),
                                                self.item_row0.data_ptr(), self.item_rows.data_ptr(), self.n_items,
                                                self.max_cols, int(self.vec4), sdt, st), "obst_opt_sm3_stage")
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
            raw = self._scratch("sums_raw", 4 * len(self.variables), torch.float64)
            self._stage(N.STAGE_REDUCE, work, flags=2, sums=raw)
            if getattr(self.params, "grad_reducer", None) is not None:
                self.params.grad_reducer.agree(raw)
        for name, args in plan.stages:
            self._run_stage(name, args, work, lr, step_count)
        if any("rezero" in v.name for v in self.variables):  # src/optimizer/__init__.py:46-47
            if not all("rezero" in v.name for v in self.variables):
                raise NotImplementedError("rezero_lr_multiplier: rezero and other variables share one generic chain")
            self._stage(N.STAGE_LEARNING_RATE, work, f=[float(self.params.rezero_lr_multiplier)])
        self._stage(N.STAGE_APPLY, work, f=[float(lr) * float(self.params.weight_decay)])
        self.wnorm_valid = False
        if refresh_transposed:
            self.refresh_transposed()

    def step(self, lr: float, refresh_transposed: bool = True, step_count: int = 1, preview: bool = False,
             chained: bool = False):
        """One update of every variable.
        preview / chained: the reference's macro-batching (src/run/train.py:35-74).  A non-final macro-batch
        (preview) moves the optimizer state on but leaves the master weights alone; the shadows receive the weights
        the next macro-batch's graph reads, var.value - update in the activation dtype.  `chained` says the graph of
        THIS macro-batch read such chained weights, so the clip norm and the weight decay use the shadows' values."""
        if self.scalar_opt is not None:
            if preview or chained:
                raise NotImplementedError("macro_batching > 1 with rank-0 (rezero) variables")
            self.scalar_opt.step(lr, False, step_count)
        if isinstance(self.plan, GenericPlan):
            if preview or chained:
                raise NotImplementedError("macro_batching > 1 needs the fused optimizer chain "
                                          "([adaptive_clip-]sm3-momentum:<b>:1:<n>-learning_rate)")
            return self._step_generic(lr, step_count, refresh_transposed)
        lib = native.load()
        st = native.stream_ptr()
        p, plan = self.params, self.plan
        nt = len(self.variables)
        act, sdt = native.dtype_code(self.act_dtype), native.dtype_code(self.state_dtype)
        # gradient hand-over dtype: 2 = stored as bf16 (flat_grad_lp), 1 = fp32 rounded to bf16 on read, 0 = fp32
        grad_bf16 = 2 if self.grad_lp else int(self.act_dtype == torch.bfloat16)
        tp, it, ir0, irs = (self.table.data_ptr(), self.item_tensor.data_ptr(), self.item_row0.data_ptr(),
                            self.item_rows.data_ptr())
        single = getattr(getattr(p, "grad_reducer", None), "world", 1) == 1
        # buckets the data-parallel reducer did not hand over (single rank: all of them, in one launch)
        self.hand_over(None, with_norms=single and plan.clip > 0)
        macro = preview or chained
        carry = 1 if (self.wnorm_valid and not macro) else 0  # the chain's weight norms are read from the shadows
        if plan.clip > 0:
            # sum g^2: taken by the hand-over on a single rank (flags 2|4: keep it, add only the weight half);
            # with several ranks the buckets were all-reduced after the hand-over, so it is taken here.
            # sum cast(p)^2: carried over from the previous update (flag 1) or, on the first step, read here.
            have_g = self.grad_lp and self.g2_keys == set(self.item_ranges)
            flags = carry | (6 if have_g else 0) | (8 if macro else 0)
            native.check(lib.obst_opt_norms(tp, nt, it, ir0, irs, self.n_items, act, grad_bf16 & 1,
                                            self.norms.data_ptr(), flags, self.wnext.data_ptr(), 0, nt, st),
                         "obst_opt_norms")
            if not single:  # float64 atomics in a per-rank order: all replicas take rank 0's sums
                p.grad_reducer.agree(self.norms)
        ut, ur0, urs = (t.data_ptr() for t in self.update_items)
        native.check(lib.obst_opt_update(tp, nt, ut, ur0, urs, self.n_update_items, self.max_cols, int(self.vec4), act, sdt,
                                         grad_bf16, self.norms.data_ptr(), self.wnext.data_ptr(), plan.clip, plan.beta,
                                         int(plan.nesterov), float(lr), float(p.weight_decay),
                                         int(preview) | (2 if macro else 0), st), "obst_opt_update")
        native.check(lib.obst_opt_finalize(tp, nt, sdt, st), "obst_opt_finalize")
        self.wnorm_valid = plan.clip > 0 and not macro
        if macro:
            self.wnext.zero_()
        if refresh_transposed:  # the update rewrote the shadows in place: transposed copies follow
            self.refresh_transposed()


def _scalar_plan(plan) -> GenericPlan:
    """The chain as it runs for a rank-0 variable: sm3 and novograd hand over to adam
    (src/optimizer/optimizers.py:48-49,61-62)."""
    if isinstance(plan, FusedPlan):
        stages = ([("adaptive_clip", (repr(plan.clip),))] if plan.clip > 0 else []) + \
                 [("adam", ()), ("momentum", (repr(plan.beta), "1", str(int(plan.nesterov)))), ("learning_rate", ())]
    else:
        stages = [("adam", ()) if n in ("sm3", "novograd") else (n, a) for n, a in plan.stages]
    return GenericPlan(tuple(stages))


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
    opt_state.step(ctx.learning_rate, step_count=ctx.step_count, preview=not ctx.assign,
                   chained=bool(getattr(params, "macro_chained", False)))
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
    """Differentiate the recorded step and run the optimizer.  Same contract as the reference
    (src/optimizer/__init__.py:69-186): returns (update_ops, learning_rate).
      fn == "update"      the committed update (`ctx.assign`)
      fn == "accumulate"  gradients are only accumulated (this implementation's micro-batches; the name is the one
                          the reference's docstring reserves for gradient accumulation)
      anything else       the reference's non-assign arm ("add" in src/run/train.py:54): the update is computed and
                          chained into the next macro-batch, the master weights stay (FusedOptimizer.step)"""
    state = _ensure_state(params)
    update_ops: typing.List[str] = []
    learning_rate_ctx = get_learning_rate(params, loss_list, update_ops)
    learning_rate = float(learning_rate_ctx.learning_rate)
    reducer = getattr(params, "grad_reducer", None)
    if reducer is not None:
        reducer.begin_backward(last=(fn != "accumulate"))  # every optimizer call needs the reduced gradient
    _backward(params)
    if fn == "accumulate":
        return update_ops, learning_rate
    if reducer is not None:
        reducer.finish()
    ctx = OptimizerCtx(params, update_ops, learning_rate, int(params.global_step) + 1, fn == "update")
    update(ctx)
    return ctx.update_ops, learning_rate
