"""Variables and the linear layer (src/model/backend.py of the reference).

* `OrthogonalInit` / `normal_var` / `orthogonal_var` / `get_var` keep the reference's init rules, its
  full-scope-name variable identity and its cross-depth `shared` rule (src/model/backend.py:18-105).
* `linear` is the reference's `einsum(tensor, orthogonal_var(old + new, fan_in=old))`
  (src/model/backend.py:108-110), lowered onto obst_gemm: dense when old/new are disjoint, grouped (one GEMM per
  head) when a dim is in both.  The weight is read where it lies for forward, dgrad and wgrad.
"""
import typing
import zlib

import numpy as np
import torch

from .. import native
from ..dataclass import BlockArgs, ModelParameter
from ..utils_core import current_scope, random_name, scope_list, scoped
from ..utils_mtf import DIM_LIST, Tensor, deduplicate, shape_size, shape_sub


class Variable:
    """One trainable tensor: fp32 master slice, activation-dtype shadow (what the graph reads), fp32 gradient."""

    def __init__(self, params: ModelParameter, full_name: str, dims: DIM_LIST, init: np.ndarray, trainable: bool):
        dev = params.device
        self.full_name = full_name
        self.name = full_name
        self.dims = list(dims)
        self.shape = self.dims
        self.trainable = trainable
        init = np.asarray(init, dtype=np.float32)  # (np.ascontiguousarray would turn a rank-0 array into rank 1)
        self.data = torch.tensor(init, dtype=torch.float32).to(dev).contiguous()  # slice_dtype fp32
        self.tril_s = 0          # >0: causal mask folded into the shadow (token-mixing maps)
        self.value: typing.Optional[torch.Tensor] = None
        self.value_t: typing.Optional[torch.Tensor] = None  # transposed shadow, see transposed()
        self._t_view: typing.Optional[typing.Tuple[int, int, int]] = None
        self.grad: typing.Optional[torch.Tensor] = None  # bound by the optimizer to a view of its flat buffer
        self._act_dtype = params.variable_dtype.activation_dtype
        # macro-batching chains would-be weights through the shadow while the master stays put
        # (src/run/train.py:63-69), so the shadow needs a buffer of its own even when the dtypes agree
        self._own_shadow = int(getattr(params, "macro_batching", 1)) > 1

    def refresh(self):
        """(Re)build the shadow: cast of the fp32 slice to the activation dtype (mtf.Variable.value)."""
        if self._act_dtype == torch.float32 and self.tril_s == 0 and not self._own_shadow:
            self.value = self.data
        else:
            if self.value is None or self.value is self.data:
                self.value = torch.empty_like(self.data, dtype=self._act_dtype)
            native.cast_param(self.data, self.value, self.tril_s)
        self.refresh_t()

    def refresh_t(self):
        """Rewrite the transposed shadow (if one was asked for) from the shadow; the optimizer calls this after
        every update, since its kernels rewrite `value` in place."""
        if self.value_t is not None:
            native.transpose_param(self.value.view(self._t_view), self.value_t)
        self.refresh_folded()

    # -- a group norm folded into this weight (basic._mid_norm_out_folded): (gamma . W)^T, gamma^T W, beta^T W ----
    folded = None  # (gamma variable, beta variable, heads, kin, kout, wg_t, cvec, dvec, gamma_f, wg)

    def fold_norm(self, gamma: typing.Optional["Variable"], beta: typing.Optional["Variable"], heads: int, kin: int,
                  kout: int):
        if self.folded is None or self.folded[0] is not gamma or self.folded[1] is not beta:
            dev = self.value.device
            self.folded = (gamma, beta, heads, kin, kout,
                           torch.empty((heads, kout, kin), dtype=self.value.dtype, device=dev),
                           torch.empty((heads, kout), dtype=torch.float32, device=dev),
                           torch.empty((heads, kout), dtype=torch.float32, device=dev),
                           torch.empty((heads, kin), dtype=torch.float32, device=dev),
                           torch.empty((heads, kin, kout), dtype=self.value.dtype, device=dev))
            self.refresh_folded()
        return self.folded[5:]

    def fold_entry(self):
        """(w, gamma, beta, wg_t, cvec, dvec, gamma_f, wg): one row of a native.fold_items table."""
        gamma, beta, heads, kin, kout, wg_t, cvec, dvec, gamma_f, wg = self.folded
        return (self.value, None if gamma is None else gamma.value, None if beta is None else beta.value, wg_t, cvec,
                dvec, gamma_f, wg)

    def refresh_folded(self):
        """Called whenever this weight or the norm's scale / shift moved outside the optimizer (creation, weight
        import, checkpoint restore); the optimizer refreshes all folded weights of a shape in one launch."""
        if self.folded is None:
            return
        heads, kin, kout = self.folded[2:5]
        native.fold_norm_weight(native.fold_items([self.fold_entry()]), 1, heads, kin, kout)

    def transposed(self, batch: int, rows: int, cols: int) -> torch.Tensor:
        """The shadow, seen as [batch, rows, cols], stored as [batch, cols, rows].  Forward GEMMs read the weight
        as their B operand with the contraction dim second-to-last (MN-major); the tcgen05 kernel is 12-17 %
        faster on a K-major B, and a weight is tiny next to an activation, so a second copy is kept instead."""
        view = (batch, rows, cols)
        if self.value_t is None or self._t_view != view:
            if batch * rows * cols != self.value.numel():
                raise ValueError(f"transposed{view} does not match {self.name} {tuple(self.value.shape)}")
            self._t_view = view
            self.value_t = torch.empty((batch, cols, rows), dtype=self.value.dtype, device=self.value.device)
            self.refresh_t()
        return self.value_t

    def set_masked(self, s: int):
        if self.tril_s != s:
            self.tril_s = s
            self.value = None
            self.refresh()

    @property
    def size(self):
        return self.data.numel()

    def ensure_grad(self):
        if self.grad is None:
            self.grad = torch.zeros_like(self.data)
        return self.grad


def _seed_for(params: ModelParameter, full_name: str) -> int:
    # the reference draws `random.randint` per variable (src/model/backend.py:24); here: stable per name
    return (zlib.crc32(full_name.encode()) + 7919 * int(params.init_seed)) % (2 ** 32)


class OrthogonalInit:
    """src/model/backend.py:18-40"""

    def __init__(self, params: ModelParameter, shape: DIM_LIST, is_last: bool, fan_in_dims=None):
        if fan_in_dims is None:
            fan_in_dims = []
        self.params = params
        self.sizes = [d.size for d in shape]
        fan_in = int(np.prod([d.size for d in fan_in_dims]))
        fan_out = int(np.prod(self.sizes)) // fan_in
        self.transpose = fan_out > fan_in
        self.shape = (fan_out, fan_in) if self.transpose else (fan_in, fan_out)
        self.scale_by_depth = params.scale_by_depth and is_last

    def __call__(self, seed: int) -> np.ndarray:
        g = torch.Generator(device="cpu").manual_seed(seed)
        a = torch.randn(self.shape, generator=g, dtype=torch.float32)
        dev = self.params.device if (torch.cuda.is_available() and a.numel() > 1 << 18) else "cpu"
        q, r = torch.linalg.qr(a.to(dev).double())  # init-time only (not on the hot path)
        q = q * torch.sign(torch.diagonal(r))
        if self.transpose:
            q = q.t()
        out = q.reshape(self.sizes)
        if self.scale_by_depth:
            out = out / self.params.depth ** 0.5
        return out.float().cpu().numpy()


class NormalInit:
    def __init__(self, sizes, stddev: float, mean: float):
        self.sizes, self.stddev, self.mean = sizes, stddev, mean

    def __call__(self, seed: int) -> np.ndarray:
        g = torch.Generator(device="cpu").manual_seed(seed)
        return (self.mean + self.stddev * torch.randn(self.sizes, generator=g, dtype=torch.float32)).numpy()


def get_variable(params: ModelParameter, name: str, shape: DIM_LIST, initializer, trainable: bool) -> Variable:
    """src/utils_mtf.py:271-280: identity = full scope name; creation order = graph order."""
    full_name = f"{current_scope()}/{name}"
    if full_name in params.variable_cache:
        return params.variable_cache[full_name]
    shape = deduplicate(list(shape))
    index = len(params.variable_cache)
    imported = getattr(params, "import_variables", None)
    restored = None
    if getattr(params, "restore_entries", None) is not None:  # obst.run.checkpoint.load_checkpoint
        from ..run.checkpoint import take_entry
        restored = take_entry(params, full_name)
    if restored is not None:
        init = restored.float().numpy().reshape([d.size for d in shape])
    elif imported is not None:
        init = np.asarray(imported[index], dtype=np.float32).reshape([d.size for d in shape])
    else:
        init = initializer(_seed_for(params, full_name))
    var = Variable(params, full_name, shape, init, trainable)
    var.refresh()
    params.variable_cache[full_name] = var
    return var


def get_var(args: BlockArgs, shape: DIM_LIST, initializer) -> Variable:
    """src/model/backend.py:43-94.  Without `shared` every call site owns a variable.  With it, the n-th shared
    request of a (block_config index, layer function) pair at depth >= 1 returns the n-th variable created at
    depth 0."""
    params = args.params

    def _var():
        return get_variable(params, random_name("get_variable"), shape, initializer, True)

    if "shared" not in args:
        return _var()

    scope = scope_list()
    body_idx = scope.index("body0v") + 1
    if scope[body_idx].startswith("recompute_grad"):
        # 'checkpoint' strategy: recompute_grad's own scope sits between body and block.  The reference parses THAT as
        # the block name and then divides by zero (src/model/backend.py:59-75): shared variables under 'checkpoint'
        # do not run there; here the block name is simply read one level deeper.
        body_idx += 1
    block, full_fn_name = scope[body_idx:body_idx + 2]
    block, config = block.split('_')
    first_block = block == '0'
    fn_name = ''.join(c for c in full_fn_name if not c.isdigit())

    cache = params.cached_parameters
    for idx in (config, fn_name):
        cache = cache.setdefault(idx, {})
    # call-site ordinal inside this (depth, config) block; reset_shared_ordinals() rewinds it for a re-run
    seen = cache.setdefault("seen", {})
    key = (block, config)
    ordinal = seen.get(key, 0)
    seen[key] = ordinal + 1
    store = cache.setdefault("vars", [])
    if first_block:
        var = _var()
        if ordinal >= len(store):
            store.append(var)
        return store[ordinal]
    return store[ordinal % len(store)]


def reset_shared_ordinals(params: ModelParameter, block: str, config: str):
    """Called before a reversible sub-block is re-executed in backward (src/model/revnet.py:59 clones the forward
    ops, i.e. reuses the same variables): the shared call sites must resolve to the same variables again."""
    for fn_cache in params.cached_parameters.get(config, {}).values():
        fn_cache.get("seen", {}).pop((block, config), None)


def orthogonal_var(args: BlockArgs, shape: DIM_LIST, fan_in_dims=None) -> Variable:
    shape = deduplicate(list(shape))
    return scoped("orthogonal_var", get_var, args, shape,
                  OrthogonalInit(args.params, shape, args.is_last, fan_in_dims))


def normal_var(args: BlockArgs, shape: DIM_LIST, stddev: float = 0.02, mean: float = 0.) -> Variable:
    shape = deduplicate(list(shape))
    return scoped("normal_var", get_var, args, shape, NormalInit([d.size for d in shape], stddev, mean))


# ----------------------------------------------------------------------------------------------------------
# tape
# ----------------------------------------------------------------------------------------------------------
class TapeEntry:
    """Backward of one executed layer.  `fn(dy, mask, acc)` returns dx.
    out_relu     : the layer's ReLU output, if it ended in one; the entry that produces this layer's dy may fold
                   [out_relu > 0] into its own epilogue (supports_mask) instead of a separate pass.
    supports_acc : fn can add its dx into an existing buffer (`acc`) instead of allocating."""

    __slots__ = ("fn", "out_relu", "supports_mask", "supports_acc")

    def __init__(self, fn, out_relu=None, supports_mask=False, supports_acc=False):
        self.fn = fn
        self.out_relu = out_relu
        self.supports_mask = supports_mask
        self.supports_acc = supports_acc


def run_tape(tape: typing.List[TapeEntry], dy: typing.Optional[torch.Tensor],
             acc: typing.Optional[torch.Tensor] = None) -> typing.Optional[torch.Tensor]:
    """Reverse sweep.  `acc`: buffer the final dx is to be ADDED to (dx2 = dy2 + J^T dy1, src/model/revnet.py:120)."""
    for i in range(len(tape) - 1, -1, -1):
        entry = tape[i]
        mask = tape[i - 1].out_relu if i > 0 else None
        last = i == 0
        use_acc = acc if (last and entry.supports_acc) else None
        if mask is not None and not entry.supports_mask:
            dy = entry.fn(dy, None, use_acc)
            native.act_bwd(0, mask, dy, dy)
        else:
            dy = entry.fn(dy, mask, use_acc)
        if last and acc is not None and use_acc is None:
            native.axpby(acc, dy, 1.0, acc)
            dy = acc
    return dy


# ----------------------------------------------------------------------------------------------------------
# linear
# ----------------------------------------------------------------------------------------------------------
def linear(args: BlockArgs, old: DIM_LIST, new: DIM_LIST, relu: bool = False,
           relu_stats: typing.Optional[torch.Tensor] = None) -> Tensor:
    """einsum([tensor, W[old + new]]) -> (tensor.shape - old) + new, optional fused ReLU.
    If `args.params.residual` is set (by the reversible driver) the result is combined with that buffer in the
    GEMM epilogue: out = residual +/- x@W, in place.
    relu_stats: fp32 [rows, N / 128, 2] buffer; with it the ReLU epilogue also emits the per-row partial sums of a
    group norm that follows (obst_gemm epilogue 7)."""
    params = args.params
    x = args.tensor
    old, new = list(old), list(new)
    var = orthogonal_var(args, old + new, old)
    group = [d for d in old if d in new]
    # MTF's einsum only contracts dims the INPUT carries: an `old` dim missing from the tensor (reduced_half_linear:
    # the head dim after sum_heads, src/model/basic.py:89-90) is simply a produced dim, and the same input feeds
    # every group
    absent = [d for d in group if d not in x.dims]
    contract = [d for d in old if d not in group and d in x.dims]
    produce = [d for d in new if d not in group]
    if len(group) > 1:
        raise NotImplementedError(f"linear: more than one grouped dim ({group}) is not supported on this path")
    present_group = [d for d in group if d not in absent]
    lead = shape_sub(x.shape, old)
    if x.dims != lead + present_group + contract:
        raise NotImplementedError(f"linear: contracted dims {old} must be the trailing dims of {x.dims}")
    out_dims = deduplicate(lead + new)
    if out_dims != lead + group + produce:
        raise NotImplementedError(f"linear: unsupported output layout {out_dims}")
    M, G, Kc, Np = shape_size(lead), shape_size(group), shape_size(contract), shape_size(produce)
    if [d.size for d in var.dims] != [d.size for d in group + [d for d in old if d not in group] + produce]:
        raise NotImplementedError(f"linear: weight layout {var.dims} is not [group, contract, produce]")
    if shape_size([d for d in old if d not in group]) != Kc:
        raise NotImplementedError(f"linear: contracted dims {old} are only partly present in {x.dims}")
    be = params.gemm_backend
    dt = x.data.dtype
    w = var.value
    wt = var.transposed(G, Kc, Np)  # [G, Np, Kc]: the forward B operand, K-major
    residual = params.residual
    xin = x.data
    xg = 0 if absent else Kc      # stride of the input along the group (0: one input broadcast to every group)
    xld = Kc if absent else G * Kc
    if residual is not None:
        params.residual = None
        buf, sign, out = residual  # out = buf +/- x@W; out is buf itself for the in-place (reversible) case
        if relu or tuple(buf.shape) != tuple(d.size for d in out_dims):
            raise NotImplementedError("linear: residual fusion needs a plain final linear of matching shape")
        native.gemm(native.operand(xin, xld, 1, s_glo=xg), native.operand(wt, Kc, 1, s_glo=Kc * Np), out, M, Np, Kc,
                    GLO=G, d_ld=G * Np, d_glo=Np, R=buf, epi=2 if sign > 0 else 3, dtype_in=dt, backend=be)
    else:
        out = torch.empty([d.size for d in out_dims], dtype=dt, device=xin.device)
        if relu_stats is not None:
            if not relu or G != 1:
                raise NotImplementedError("linear: the statistics epilogue rides on an ungrouped ReLU linear")
            native.gemm(native.operand(xin, xld, 1), native.operand(wt, Kc, 1), out, M, Np, Kc, d_ld=Np, epi=7,
                        dtype_in=dt, backend=2, aux=(relu_stats,))
        else:
            native.gemm(native.operand(xin, xld, 1, s_glo=xg), native.operand(wt, Kc, 1, s_glo=Kc * Np), out, M, Np,
                        Kc, GLO=G, d_ld=G * Np, d_glo=Np, epi=1 if relu else 0, dtype_in=dt, backend=be)

    if params.tape is not None:
        def bwd(dy: torch.Tensor, mask, acc):
            # wgrad: dW[g,kc,np] += sum_m x[m,g,kc] dy[m,g,np]   (both operands read MN-major, fp32 atomics)
            native.gemm(native.operand(xin, 1, xld, s_glo=xg), native.operand(dy, 1, G * Np, s_glo=Np),
                        var.ensure_grad(), Kc, Np, M, GLO=G, d_ld=Np, d_glo=Kc * Np, accumulate=True, dtype_in=dt,
                        backend=be)
            dx = torch.empty_like(xin)
            if absent:
                # dx[m,kc] = sum_g sum_np dy[m,g,np] W[g,kc,np]: the group is an outer reduction index
                native.gemm(native.operand(dy, G * Np, 1, s_ko=Np), native.operand(w, Np, 1, s_ko=Kc * Np), dx, M, Kc,
                            Np, KO=G, d_ld=Kc, R=mask, epi=4 if mask is not None else 0, dtype_in=dt, backend=be)
                return dx
            # dgrad: dx[m,g,kc] = sum_np dy[m,g,np] W[g,kc,np]   (+ ReLU mask of the producer folded in)
            native.gemm(native.operand(dy, G * Np, 1, s_glo=Np), native.operand(w, Np, 1, s_glo=Kc * Np), dx, M, Kc,
                        Np, GLO=G, d_ld=G * Kc, d_glo=Kc, R=mask, epi=4 if mask is not None else 0, dtype_in=dt,
                        backend=be)
            return dx

        params.tape.append(TapeEntry(bwd, out_relu=out if relu else None, supports_mask=True))
    return Tensor(out, out_dims)


def linear_to_features(args: BlockArgs, old: typing.Optional[DIM_LIST] = None) -> Tensor:
    return linear(args, old, args.params.feature_dims)


def linear_from_features(args: BlockArgs, new: typing.Optional[DIM_LIST] = None) -> Tensor:
    return linear(args, args.params.feature_dims, new)
