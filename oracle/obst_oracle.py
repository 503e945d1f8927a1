"""CPU oracle: a NumPy restatement of the OBST mixer training step.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker / reported CPU baseline.  The product (homebrewnlp-mtf_b200/) never does.

PARITY: pinned against THE REFERENCE'S OWN PYTHON, EXECUTED.  The reference holds no golden vectors for this path
(its tests assert statistics of initialisers only) and its two tensor libraries cannot be
installed here (tensorflow==2.7.0 + mesh-tensorflow: requirements.txt:2-3, Python 3.12, no network), but the
reference itself is pure Python: tests/golden/make_reference_vectors.py imports /root/reference's src.* UNMODIFIED
on top of eager NumPy stand-ins for those two libraries (oracle/ref_shim: named-dimension arithmetic, einsum,
reductions and every op's gradient rule as mesh-tensorflow defines them; TensorFlow functions as their NumPy
namesakes) and records what src/run/train.py:get_train_model -- build(), the RevGradOp reverse sweep with its
reconstruction, get_optimizer with every stage -- returns.  tests/test_oracle_pins.py holds this file to those records
(tests/golden/reference_*.npz): logits, loss, accuracy, every gradient, every updated weight and optimizer slot over
consecutive steps, eight optimizer strings, the macro-batching chain.  What the stand-ins cannot pin is XLA's
arithmetic itself (summation order, bf16 rounding points): those stay with (a) the line-by-line citations below,
(b) the reference-authored statistical invariants of tests/test_oracle.py, (c) the torch-autograd cross-check of
every hand-written backward formula.

Citations are paths under /root/reference.
"""
import json
import math
import typing

import numpy as np

GN_EPS = 1e-5  # src/model/normalization.py:28


# ----------------------------------------------------------------------------------------------------------
# configuration: the derived sizes of src/dataclass.py:184-342 for the language ("gpt") mode
# ----------------------------------------------------------------------------------------------------------
class Shapes:
    """Sizes derived exactly as ModelParameter does (src/dataclass.py:221-234,283-285; src/model/__init__.py:76-78)."""

    def __init__(self, config: typing.Dict[str, typing.Any]):
        c = dict(config)
        self.config = c
        self.heads = int(c.get("heads", 8))
        self.features_per_head = c.get("features_per_head")
        features = c.get("features")
        if features is None and self.features_per_head is None:
            raise ValueError("Either features or features_per_head has to be specified")  # src/dataclass.py:221
        if features is None:
            features = self.features_per_head * self.heads
        if self.features_per_head is None:
            self.features_per_head = features // self.heads
        self.features = int(features)
        self.sequence_length = int(c.get("sequence_length", 32))
        self.train_batch_size = int(c.get("train_batch_size", 1))
        self.depth = int(c.get("depth", 16))
        self.vocab_size = int(c.get("vocab_size", 256))
        self.group_linear_factor = c.get("group_linear_factor", 2)
        mm = c.get("intermediate_feed_forward_multiplier_multiplier")
        m = c.get("intermediate_feed_forward_multiplier")
        if mm is not None:  # src/dataclass.py:230-232
            m = self.group_linear_factor * mm / self.heads
        if m is None:
            m = self.group_linear_factor / self.heads
        self.intermediate = int(self.heads * self.features_per_head * m)  # src/dataclass.py:283-285
        self.vocab_weight_factorization = c.get("vocab_weight_factorization", 0.125)
        self.intermediate0 = int(self.intermediate * self.vocab_weight_factorization)  # src/model/__init__.py:78
        self.group_features = int(self.features_per_head * self.group_linear_factor)  # src/utils_mtf.py:376-380
        self.embedding_stddev = c.get("embedding_stddev", 0.04)
        self.z_loss = c.get("z_loss", 1e-4)
        self.learning_rate = c.get("learning_rate", 5e-5)
        self.learning_rate_config = c.get("learning_rate_config", {})
        self.weight_decay = c.get("weight_decay", 0.001)
        self.optimizer = c.get("optimizer", "learning_rate")
        self.calculation_dtype = c.get("calculation_dtype", "float32")
        self.optimizer_slice_dtype = c.get("optimizer_slice_dtype", "float32")
        self.memory_reduction_strategy = c.get("memory_reduction_strategy", "revnet")
        # `skip` of each block_config entry: out += block_input under the 'none' / 'checkpoint' strategies
        # (src/model/frontend.py:29-30; src/dataclass.py BlockConfig default False)
        self.skip = [bool(b.get("skip", False)) for b in c.get("block_config", [{}, {}])]

    # short names used in the math below
    @property
    def dims(self):
        return (self.train_batch_size, self.sequence_length, self.heads, self.features_per_head, self.features,
                self.intermediate, self.group_features, self.intermediate0, self.vocab_size, self.depth)


def load_config(path: str) -> typing.Dict[str, typing.Any]:
    with open(path) as f:
        return json.load(f)


TINY_CONFIG = {  # BASELINE.json configs[0]: "tiny mixer LM (2 layers, d_model=256, ctx=256, batch 8) fp32"
    "embedding_stddev": 0.004, "calc_accuracy": True, "scale_by_depth": True,
    "block_config": [
        {"layer": ["norm-shift-scale-features-group",
                   "bottleneck_group_linear-in:relu-mid:relu-mid:norm-mid:shift-mid:scale-mid:features"]},
        {"layer": ["norm-shift-scale-features-group",
                   "attention-biased_attention_map-absolute-input_as_value-shared",
                   "norm-shift-scale-features-group", "activation-gelu",
                   "attention-biased_attention_map-absolute-input_as_value-shared"]}],
    "group_linear_factor": 2, "intermediate_feed_forward_multiplier_multiplier": 0.5, "depth": 2,
    "use_initial_position_embedding": False, "sequence_length": 256, "features_per_head": 32, "heads": 8,
    "train_batch_size": 8, "data_seed": 134567, "vocab_size": 256, "model_mode": "gpt", "use_language": True,
    "learning_rate": 0.01, "opt_beta1": 0.9, "memory_reduction_strategy": "revnet", "opt_beta2": 0.99,
    "optimizer": "adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", "weight_decay": 0.0001,
    "macro_batching": 1, "calculation_dtype": "float32", "storage_dtype": "float32",
    "optimizer_slice_dtype": "float32", "slice_dtype": "float32",
    "learning_rate_config": {"linear_warmup": {"final_step": 4096}},
}


# ----------------------------------------------------------------------------------------------------------
# initialisation (src/model/backend.py:18-40,97-105; src/model/embedding.py:175-178)
# ----------------------------------------------------------------------------------------------------------
def orthogonal_init(rng: np.random.Generator, sizes: typing.Sequence[int], fan_in: int) -> np.ndarray:
    """OrthogonalInit.__call__ (src/model/backend.py:28-40): QR of a normal [max,min] matrix, sign-fixed,
    transposed when fan_out > fan_in, reshaped to the variable's dims.  scale_by_depth is never applied on
    this path (BlockArgs.__call__ drops is_last, src/dataclass.py:396)."""
    total = int(np.prod(sizes))
    fan_out = total // fan_in
    transpose = fan_out > fan_in
    shape = (fan_out, fan_in) if transpose else (fan_in, fan_out)
    q, r = np.linalg.qr(rng.standard_normal(shape))
    q = q * np.sign(np.diag(r))
    if transpose:
        q = q.T
    return np.ascontiguousarray(q.reshape(sizes)).astype(np.float32)


def param_names(sh: Shapes) -> typing.List[str]:
    """Creation order of the trainable variables (depth-major, as the graph builds them)."""
    names = ["embed", "w0"]
    for i in range(sh.depth):
        names += [f"g0_{i}", f"b0_{i}", f"win_{i}", f"wmid_{i}", f"gm_{i}", f"bm_{i}", f"wout_{i}",
                  f"g1_{i}", f"b1_{i}"]
        if i == 0:
            names.append("mix_a")
        names += [f"g2_{i}", f"b2_{i}"]
        if i == 0:
            names.append("mix_b")
    names.append("wo")
    return names


def init_params(sh: Shapes, seed: int = 0) -> typing.Dict[str, np.ndarray]:
    """Seeded draw following the reference's init rules.  (The reference seeds each OrthogonalInit with
    Python's global `random`, src/model/backend.py:24, so parity runs IMPORT these weights on both sides.)"""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    rng = np.random.default_rng(seed)
    p: typing.Dict[str, np.ndarray] = {}

    def normal(shape, mean, std):  # normal_var, src/model/backend.py:103-105
        return (mean + std * rng.standard_normal(shape)).astype(np.float32)

    p["embed"] = normal((V, I0), 0.0, sh.embedding_stddev)  # src/model/embedding.py:175-178,230
    p["w0"] = orthogonal_init(rng, (1, I0, H, K), fan_in=I0)  # src/model/__init__.py:82
    for i in range(L):
        p[f"g0_{i}"] = normal((H, K), 1.0, 0.02)  # src/model/normalization.py:30-33
        p[f"b0_{i}"] = normal((H, K), 0.0, 0.02)
        p[f"win_{i}"] = orthogonal_init(rng, (H, K, I), fan_in=H * K)  # src/model/basic.py:33,124
        p[f"wmid_{i}"] = orthogonal_init(rng, (I, H, K2), fan_in=I)
        p[f"gm_{i}"] = normal((H, K2), 1.0, 0.02)
        p[f"bm_{i}"] = normal((H, K2), 0.0, 0.02)
        p[f"wout_{i}"] = orthogonal_init(rng, (H, K2, K), fan_in=H * K2)
        p[f"g1_{i}"] = normal((H, K), 1.0, 0.02)
        p[f"b1_{i}"] = normal((H, K), 0.0, 0.02)
        if i == 0:  # "shared": one variable per call site, reused at every depth (src/model/backend.py:50-94)
            p["mix_a"] = normal((H, S, S), 0.0, sh.embedding_stddev)  # src/model/spatial.py:19-23
        p[f"g2_{i}"] = normal((H, K), 1.0, 0.02)
        p[f"b2_{i}"] = normal((H, K), 0.0, 0.02)
        if i == 0:
            p["mix_b"] = normal((H, S, S), 0.0, sh.embedding_stddev)
    # output_embedding = "absolute-orthogonal": orthogonal_var WITHOUT fan_in dims -> fan_in 1 -> QR of a
    # column = one unit-norm vector (src/dataclass.py:41, src/model/backend.py:20-29)
    p["wo"] = orthogonal_init(rng, (H, K, 1, V), fan_in=1)
    assert list(p.keys()) == param_names(sh)
    return p


def init_params_fast(sh: Shapes, seed: int = 0) -> typing.Dict[str, np.ndarray]:
    """TIMING ONLY (bench.py's CPU baseline): same tensors and scales as init_params, but the orthogonal weights
    are drawn as N(0, 1/fan) instead of through a QR (which alone takes ~20 s for the 32big parameter set)."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    rng = np.random.default_rng(seed)
    p: typing.Dict[str, np.ndarray] = {}

    def normal(shape, mean, std):
        return (mean + std * rng.standard_normal(shape, dtype=np.float32)).astype(np.float32)

    p["embed"] = normal((V, I0), 0.0, sh.embedding_stddev)
    p["w0"] = normal((1, I0, H, K), 0.0, (1.0 / max(I0, F)) ** 0.5)
    for i in range(L):
        p[f"g0_{i}"] = normal((H, K), 1.0, 0.02)
        p[f"b0_{i}"] = normal((H, K), 0.0, 0.02)
        p[f"win_{i}"] = normal((H, K, I), 0.0, (1.0 / max(F, I)) ** 0.5)
        p[f"wmid_{i}"] = normal((I, H, K2), 0.0, (1.0 / max(I, H * K2)) ** 0.5)
        p[f"gm_{i}"] = normal((H, K2), 1.0, 0.02)
        p[f"bm_{i}"] = normal((H, K2), 0.0, 0.02)
        p[f"wout_{i}"] = normal((H, K2, K), 0.0, (1.0 / max(H * K2, K)) ** 0.5)
        p[f"g1_{i}"] = normal((H, K), 1.0, 0.02)
        p[f"b1_{i}"] = normal((H, K), 0.0, 0.02)
        if i == 0:
            p["mix_a"] = normal((H, S, S), 0.0, sh.embedding_stddev)
        p[f"g2_{i}"] = normal((H, K), 1.0, 0.02)
        p[f"b2_{i}"] = normal((H, K), 0.0, 0.02)
        if i == 0:
            p["mix_b"] = normal((H, S, S), 0.0, sh.embedding_stddev)
    p["wo"] = normal((H, K, 1, V), 0.0, (1.0 / (F * V)) ** 0.5)
    return p


def is_large_tensor(name: str) -> bool:
    """Weight-decay predicate of src/optimizer/__init__.py:49-57 evaluated on this model's variables:
    only the three feed-forward weights pass (norm/embed/input/output scopes are excluded)."""
    return name.startswith(("win_", "wmid_", "wout_"))


def synthetic_tokens(sh: Shapes, seed: typing.Optional[int] = None, batch: typing.Optional[int] = None):
    """x = w[:, :S], y = w[:, 1:S+1] (src/inputs.py:543-551, output_offset=1 src/dataclass.py:48)."""
    seed = sh.config.get("data_seed", 456772) if seed is None else seed
    b = sh.train_batch_size if batch is None else batch
    rng = np.random.default_rng(seed)
    w = rng.integers(0, sh.vocab_size, size=(b, sh.sequence_length + 1), dtype=np.int64).astype(np.int32)
    return np.ascontiguousarray(w[:, :-1]), np.ascontiguousarray(w[:, 1:])


# ----------------------------------------------------------------------------------------------------------
# layer math, forward and explicit backward
# ----------------------------------------------------------------------------------------------------------
def gn_fwd(x, gamma, beta):
    """norm(): src/model/normalization.py:22-34 with extras shift,scale,features,group -> per (token, head)."""
    xh = x - x.mean(-1, keepdims=True)
    r = 1.0 / np.sqrt((xh * xh).mean(-1, keepdims=True) + GN_EPS)
    z = xh * r
    return z * gamma + beta, (z, r)


def gn_bwd(dy, cache, gamma):
    z, r = cache
    red = tuple(range(dy.ndim - 2))
    dgamma = (dy * z).sum(red)
    dbeta = dy.sum(red)
    g = dy * gamma
    dx = r * (g - g.mean(-1, keepdims=True) - z * (g * z).mean(-1, keepdims=True))
    return dx, dgamma, dbeta


_C0 = 0.044715
_C1 = math.sqrt(2.0 / math.pi)


def gelu_fwd(x):
    """_gelu: src/model/activation.py:158-161 -- the cubic term is NOT scaled by sqrt(2/pi)."""
    return 0.5 * x * (1.0 + np.tanh(_C0 * x * x * x + _C1 * x))


def gelu_grad(x):
    th = np.tanh(_C0 * x * x * x + _C1 * x)
    return 0.5 * (1.0 + th) + 0.5 * x * (1.0 - th * th) * (3.0 * _C0 * x * x + _C1)


def mix_fwd(m_tril, x):
    """attention() with biased_attention_map + input_as_value: out[b,s,h,k] = sum_t M[h,s,t] tril[s,t] x[b,t,h,k]
    (src/model/spatial.py:19-23,73,77,81; mask = greater_equal(range(seq), range(_seq)), src/utils_mtf.py:411-415)."""
    # [h,s,t] @ [b,h,t,k] -> [b,h,s,k]   (contiguous copies so that BLAS takes the batched product)
    xt = np.ascontiguousarray(x.transpose(0, 2, 1, 3))
    return np.ascontiguousarray(np.matmul(m_tril[None], xt).transpose(0, 2, 1, 3))


def mix_bwd(m_tril, x, do):
    dot = np.ascontiguousarray(do.transpose(0, 2, 1, 3))  # [b,h,s,k]
    mt = np.ascontiguousarray(m_tril.transpose(0, 2, 1))
    dx = np.ascontiguousarray(np.matmul(mt[None], dot).transpose(0, 2, 1, 3))
    # dM[h,s,t] = tril * sum_{b,k} do[b,s,h,k] x[b,t,h,k]:  per head [s, b*k] @ [b*k, t]
    B, S, H, K = x.shape
    a = np.ascontiguousarray(do.transpose(2, 1, 0, 3)).reshape(H, S, B * K)
    b = np.ascontiguousarray(x.transpose(2, 0, 3, 1)).reshape(H, B * K, S)
    dm = np.matmul(a, b)
    return dx, np.tril(dm)


def cumsum_fwd(x, axis: int = 1, mean: bool = False):
    """cumsum / cummean (src/model/spatial.py:26-39): tf.cumsum along the attention dim; cummean divides by
    1 + range(dim)."""
    out = np.cumsum(x, axis=axis)
    if mean:
        shape = [1] * x.ndim
        shape[axis] = -1
        out = out / (1.0 + np.arange(x.shape[axis], dtype=x.dtype)).reshape(shape)
    return out


def cumsum_bwd(dy, axis: int = 1, mean: bool = False):
    """_cumsum_grad (src/model/spatial.py:26-28): tf.cumsum(dy, reverse=True); for cummean the division is
    differentiated first."""
    if mean:
        shape = [1] * dy.ndim
        shape[axis] = -1
        dy = dy / (1.0 + np.arange(dy.shape[axis], dtype=dy.dtype)).reshape(shape)
    return np.flip(np.cumsum(np.flip(dy, axis), axis=axis), axis)


def _ident(x):
    return x


class Block0:
    """block_config[0] = norm ; bottleneck_group_linear (src/model/basic.py:47-57,122-126)."""

    @staticmethod
    def fwd(p, i, x, keep=False, rnd=_ident):
        B, S, H, K = x.shape
        a, c0 = gn_fwd(x, p[f"g0_{i}"], p[f"b0_{i}"])
        a = rnd(a)
        win = p[f"win_{i}"]
        wmid = p[f"wmid_{i}"]
        wout = p[f"wout_{i}"]
        I = win.shape[-1]
        K2 = wmid.shape[-1]
        a2 = a.reshape(B * S, H * K)
        h1 = rnd(np.maximum(a2 @ win.reshape(H * K, I), 0))  # in: relu
        h2 = rnd(np.maximum(h1 @ wmid.reshape(I, H * K2), 0)).reshape(B, S, H, K2)  # mid: relu
        n, cm = gn_fwd(h2, p[f"gm_{i}"], p[f"bm_{i}"])  # mid: norm-shift-scale (group width K2)
        n = rnd(n)
        nh = np.ascontiguousarray(n.transpose(2, 0, 1, 3)).reshape(H, B * S, K2)
        f = np.ascontiguousarray(np.matmul(nh, wout).reshape(H, B, S, -1).transpose(1, 2, 0, 3))  # out: grouped
        cache = (a2, c0, h1, h2, n, cm) if keep else None
        return f, cache

    @staticmethod
    def bwd(p, i, cache, df, rnd=_ident):
        a2, c0, h1, h2, n, cm = cache
        B, S, H, K2 = h2.shape
        win = p[f"win_{i}"]
        wmid = p[f"wmid_{i}"]
        wout = p[f"wout_{i}"]
        I = win.shape[-1]
        K = wout.shape[-1]
        g = {}
        nh = np.ascontiguousarray(n.transpose(2, 3, 0, 1)).reshape(H, K2, B * S)
        dfh = np.ascontiguousarray(df.transpose(2, 0, 1, 3)).reshape(H, B * S, K)
        g[f"wout_{i}"] = np.matmul(nh, dfh)
        dn = np.matmul(dfh, np.ascontiguousarray(wout.transpose(0, 2, 1))).reshape(H, B, S, K2)
        dn = rnd(np.ascontiguousarray(dn.transpose(1, 2, 0, 3)))
        dh2, g[f"gm_{i}"], g[f"bm_{i}"] = gn_bwd(dn, cm, p[f"gm_{i}"])
        dh2 = rnd(dh2 * (h2 > 0)).reshape(B * S, H * K2)
        g[f"wmid_{i}"] = (h1.T @ dh2).reshape(I, H, K2)
        dh1 = rnd((dh2 @ wmid.reshape(I, H * K2).T) * (h1 > 0))
        g[f"win_{i}"] = (a2.T @ dh1).reshape(H, K, I)
        da = rnd(dh1 @ win.reshape(H * K, I).T).reshape(B, S, H, K)
        dx, g[f"g0_{i}"], g[f"b0_{i}"] = gn_bwd(da, c0, p[f"g0_{i}"])
        return dx, g


class Block1:
    """block_config[1] = norm ; mix_A ; norm ; gelu ; mix_B  (configs/32big_mixer.json:13-19)."""

    @staticmethod
    def fwd(p, i, x, keep=False, rnd=_ident):
        ma = np.tril(p["mix_a"])
        mb = np.tril(p["mix_b"])
        a, c1 = gn_fwd(x, p[f"g1_{i}"], p[f"b1_{i}"])
        a = rnd(a)
        m1 = rnd(mix_fwd(ma, a))
        y2, c2 = gn_fwd(m1, p[f"g2_{i}"], p[f"b2_{i}"])
        c = rnd(gelu_fwd(y2))
        f = mix_fwd(mb, c)
        cache = (a, c1, c2, y2, c, ma, mb) if keep else None
        return f, cache

    @staticmethod
    def bwd(p, i, cache, df, rnd=_ident):
        a, c1, c2, y2, c, ma, mb = cache
        g = {}
        dc, g["mix_b"] = mix_bwd(mb, c, df)
        dy2 = rnd(dc) * gelu_grad(y2)
        dm1, g[f"g2_{i}"], g[f"b2_{i}"] = gn_bwd(dy2, c2, p[f"g2_{i}"])
        dm1 = rnd(dm1)
        da, g["mix_a"] = mix_bwd(ma, a, dm1)
        da = rnd(da)
        dx, g[f"g1_{i}"], g[f"b1_{i}"] = gn_bwd(da, c1, p[f"g1_{i}"])
        return dx, g


BLOCKS = (Block0, Block1)


def _cast_params(params, dtype):
    return {k: v.astype(dtype) for k, v in params.items()}


def forward(sh: Shapes, params, tok_x, tok_y, dtype=np.float64, n_global: typing.Optional[int] = None,
            trace: typing.Optional[dict] = None, rnd=_ident):
    if sh.memory_reduction_strategy in ("none", "checkpoint"):
        return _forward_plain(sh, params, tok_x, tok_y, dtype, n_global, trace, rnd)
    return _forward_revnet(sh, params, tok_x, tok_y, dtype, n_global, trace, rnd)


def _head_and_loss(sh: Shapes, p, out, tok_y, Bn, n, rnd, trace):
    """_output + softmax cross-entropy with z-loss + accuracy (src/model/__init__.py:138-145,183-187;
    src/mtf_wrapper.py:64-71)."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    logits = rnd(out.reshape(Bn * S, F) @ p["wo"].reshape(F, V))
    mx = logits.max(-1, keepdims=True)
    logz = np.log(np.exp(logits - mx).sum(-1)) + mx[:, 0]
    tgt = tok_y.reshape(-1).astype(np.int64)
    nll = (logz - logits[np.arange(Bn * S), tgt]).sum()
    z2 = (logz * logz).sum()
    loss = nll / n + sh.z_loss * z2 / n
    acc = float((logits.argmax(-1) == tgt).sum()) / n
    if trace is not None:
        trace["out"] = out.copy()
        trace["logits"] = logits.copy()
        trace["logz"] = logz.copy()
        trace["loss_nll"] = nll / n
    return float(loss), acc, logits, logz, tgt


def _forward_plain(sh: Shapes, params, tok_x, tok_y, dtype, n_global, trace, rnd):
    """memory_reduction_strategy 'none' / 'checkpoint' (src/model/__init__.py:107-123): the sub-blocks are composed,
    x <- f(x) (+ x when the block_config entry sets `skip`, src/model/frontend.py:29-30).  'checkpoint' differs only
    in WHEN the intermediates are computed (mtf recompute_grad), not in any value."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    Bn = tok_x.shape[0]
    p = _cast_params(params, dtype)
    n = float(Bn * S if n_global is None else n_global)
    e = p["embed"][tok_x]
    x = rnd((e.reshape(Bn * S, I0) @ p["w0"].reshape(I0, F)).reshape(Bn, S, H, K))
    caches = []
    for i in range(L):
        for c, blk in enumerate(BLOCKS):
            f, cache = blk.fwd(p, i, x, keep=True, rnd=rnd)
            caches.append(cache)
            x = rnd(f + x) if sh.skip[c] else rnd(f)
            if trace is not None:
                trace[f"x_{i}_{c}"] = x.copy()
    loss, acc, logits, logz, tgt = _head_and_loss(sh, p, x, tok_y, Bn, n, rnd, trace)
    state = dict(p=p, e=e, out=x, logits=logits, logz=logz, tgt=tgt, n=n, tok_x=tok_x, rnd=rnd, caches=caches)
    return loss, acc, state


def _forward_revnet(sh: Shapes, params, tok_x, tok_y, dtype=np.float64, n_global: typing.Optional[int] = None,
                    trace: typing.Optional[dict] = None, rnd=_ident):
    """_input -> _body (revnet) -> _output -> _loss  (src/model/__init__.py:75-85,101-129,138-145,183-187).
    Returns (loss, accuracy, state) where state holds what backward() needs.  `n_global` = targets.size of the
    GLOBAL batch (src/mtf_wrapper.py:67: the loss normaliser), default = this batch.
    `rnd` (default identity) is applied wherever an activation tensor is materialised; passing `round_bf16` gives
    the arithmetic twin of a calculation_dtype=bfloat16 run, used by the tests to size the error budget."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    Bn = tok_x.shape[0]
    p = _cast_params(params, dtype)
    n = float(Bn * S if n_global is None else n_global)
    e = p["embed"][tok_x]  # Gather, src/model/embedding.py:118-122
    src = rnd((e.reshape(Bn * S, I0) @ p["w0"].reshape(I0, F)).reshape(Bn, S, H, K))
    x1 = src
    x2 = src
    if trace is not None:
        trace["src"] = src.copy()
    for i in range(L):
        for c, blk in enumerate(BLOCKS):
            f, _ = blk.fwd(p, i, x2, rnd=rnd)
            x1, x2 = x2, rnd(x1 + f)  # RevGradOp: y1 = x1 + f(x2); outputs (x2, y1)  src/model/revnet.py:23-24
            if trace is not None:
                trace[f"x1_{i}_{c}"] = x1.copy()
                trace[f"x2_{i}_{c}"] = x2.copy()
    out = rnd(x1 + x2)  # src/model/__init__.py:129
    logits = rnd(out.reshape(Bn * S, F) @ p["wo"].reshape(F, V))
    mx = logits.max(-1, keepdims=True)
    logz = np.log(np.exp(logits - mx).sum(-1)) + mx[:, 0]
    tgt = tok_y.reshape(-1).astype(np.int64)
    nll = (logz - logits[np.arange(Bn * S), tgt]).sum()
    z2 = (logz * logz).sum()
    loss = nll / n + sh.z_loss * z2 / n
    acc = float((logits.argmax(-1) == tgt).sum()) / n
    if trace is not None:
        trace["out"] = out.copy()
        trace["logits"] = logits.copy()
        trace["logz"] = logz.copy()
        trace["loss_nll"] = nll / n
    state = dict(p=p, e=e, x1=x1, x2=x2, out=out, logits=logits, logz=logz, tgt=tgt, n=n, tok_x=tok_x, rnd=rnd)
    return float(loss), acc, state


def backward(sh: Shapes, state, trace: typing.Optional[dict] = None):
    """Reverse sweep of src/optimizer/__init__.py:143-174 with RevGradOp.gradient (src/model/revnet.py:55-120):
    inputs of each reversible sub-block are RECONSTRUCTED (x1 = y1 - f(x2)), f is recomputed, then backpropagated."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    p = state["p"]
    logits, logz, tgt, n = state["logits"], state["logz"], state["tgt"], state["n"]
    Bn = state["tok_x"].shape[0]
    sm = np.exp(logits - logz[:, None])
    dlogits = sm / n
    dlogits[np.arange(Bn * S), tgt] -= 1.0 / n
    dlogits += (2.0 * sh.z_loss / n) * logz[:, None] * sm
    rnd = state.get("rnd", _ident)
    dlogits = rnd(dlogits)
    grads: typing.Dict[str, np.ndarray] = {}
    out2 = state["out"].reshape(Bn * S, F)
    grads["wo"] = (out2.T @ dlogits).reshape(H, K, 1, V)
    dout = rnd(dlogits @ p["wo"].reshape(F, V).T).reshape(Bn, S, H, K)
    if "caches" in state:  # 'none' / 'checkpoint': plain chain rule through the stored (or recomputed) sub-blocks
        dx = dout
        caches = state["caches"]
        for i in reversed(range(L)):
            for c in reversed(range(len(BLOCKS))):
                dxf, g = BLOCKS[c].bwd(p, i, caches[i * len(BLOCKS) + c], dx, rnd=rnd)
                for k, v in g.items():
                    grads[k] = grads[k] + v if k in grads else v
                dx = rnd(dxf + dx) if sh.skip[c] else rnd(dxf)
        dsrc = dx.reshape(Bn * S, F)
        e2 = state["e"].reshape(Bn * S, I0)
        grads["w0"] = (e2.T @ dsrc).reshape(1, I0, H, K)
        de = rnd(dsrc @ p["w0"].reshape(I0, F).T)
        dE = np.zeros((V, I0), dtype=de.dtype)
        np.add.at(dE, state["tok_x"].reshape(-1), de)
        grads["embed"] = dE
        return grads
    pp, q = state["x1"], state["x2"]  # final (x1, x2)
    dp, dq = dout, dout.copy()
    for i in reversed(range(L)):
        for c in reversed(range(len(BLOCKS))):
            blk = BLOCKS[c]
            x2_in = pp  # src/model/revnet.py:58,90
            f, cache = blk.fwd(p, i, x2_in, keep=True, rnd=rnd)  # recompute (cloned forward ops) :59-60
            x1_in = rnd(q - f)  # :89
            dx2_f, g = blk.bwd(p, i, cache, dq, rnd=rnd)
            for k, v in g.items():
                grads[k] = grads[k] + v if k in grads else v  # shared mixers accumulate over all depths
            dx1_in = dq  # :88
            dx2_in = rnd(dp + dx2_f)  # :120
            pp, q, dp, dq = x1_in, x2_in, dx1_in, dx2_in
            if trace is not None:
                trace[f"rx1_{i}_{c}"] = pp.copy()
                trace[f"rx2_{i}_{c}"] = q.copy()
    dsrc = rnd(dp + dq).reshape(Bn * S, F)
    e2 = state["e"].reshape(Bn * S, I0)
    grads["w0"] = (e2.T @ dsrc).reshape(1, I0, H, K)
    de = rnd(dsrc @ p["w0"].reshape(I0, F).T)
    dE = np.zeros((V, I0), dtype=de.dtype)
    np.add.at(dE, state["tok_x"].reshape(-1), de)  # ScatterAdd, src/model/embedding.py:39-83
    grads["embed"] = dE
    return grads


# ----------------------------------------------------------------------------------------------------------
# optimizer  (src/optimizer/__init__.py:31-66; src/optimizer/optimizers.py:13-14,60-84,114-128)
# ----------------------------------------------------------------------------------------------------------
def round_bf16(x: np.ndarray) -> np.ndarray:
    """fp32 -> bfloat16 -> fp32, round-to-nearest-even (what a cast to a bfloat16 slice does)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    out = ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)
    return np.where(np.isnan(x), x, out)


def learning_rate(sh: Shapes, global_step: int) -> float:
    """get_learning_rate + MODULES (src/optimizer/learning_rate.py:27-72); global_step is read BEFORE the
    increment, so lr == 0 on the very first step with linear_warmup."""
    lr = np.float32(sh.learning_rate)
    gs = np.float32(global_step)
    for name, cfg in sh.learning_rate_config.items():
        start = np.float32(cfg.get("start_step", 0))
        final = np.float32(cfg.get("final_step", 0))
        factor = np.float32(cfg.get("factor", 1.0))
        if name == "linear_warmup":
            warm = np.float32(gs < final)
            lr = lr * (gs / final * warm + np.float32(1) * (np.float32(1) - warm))
        elif name == "exponential_decay":
            lr = lr * np.power(factor, np.maximum(gs - start, np.float32(0)))
        elif name == "linear_decay":
            lr = lr * np.clip(np.float32(1) - (gs - start) / (final - start), 0, 1)
        elif name == "lower_bound":
            lr = np.maximum(lr, factor)
        elif name == "upper_bound":
            lr = np.minimum(lr, factor)
        else:
            raise KeyError(name)
    return float(np.float32(lr))


def init_opt_state(params) -> typing.Dict[str, typing.Dict[str, np.ndarray]]:
    """zeros, one buffer per dim + momentum (src/optimizer/backend.py:23-25; optimizers.py:64-68,123)."""
    st = {}
    for k, v in params.items():
        st[k] = {f"dim{i}": np.zeros(d, np.float32) for i, d in enumerate(v.shape)}
        st[k]["momentum"] = np.zeros(v.shape, np.float32)
    return st


def _opt_rsqrt(x):
    """opt_rsqrt, src/optimizer/optimizers.py:13-14: 1 / max(sqrt(x), 1e-5)."""
    return np.float32(1) / np.maximum(np.sqrt(x), np.float32(1e-5))


def _sum_f32(x) -> np.float32:
    return np.float32(np.sum(x.astype(np.float64)))


def _apply_stage(opt: str, args, g, w, st, hyper):
    """One registered optimizer stage (src/optimizer/optimizers.py:32-151) on one variable.  g: fp32 gradient as
    the previous stage left it; w: the weight as the graph sees it (activation-dtype cast); st: the variable's
    slots (created on first use, zeros: src/optimizer/backend.py:23-25); hyper: lr, step_count, beta1, beta2,
    global_sq (sum of squares of ALL raw gradients, for global_l2norm_clip), keep (slot rounding)."""
    f32 = np.float32
    keep = hyper["keep"]
    b1, b2, t = f32(hyper["beta1"]), f32(hyper["beta2"]), f32(hyper["step_count"])

    def slot(name, shape):
        if name not in st:
            st[name] = np.zeros(shape, f32)
        return st[name]

    def debias_momentum(beta):  # optimizers.py:17-18
        return f32(1) / (f32(1) - np.power(beta, t))

    def adam():  # optimizers.py:32-43
        p2 = slot("exp_avg_p2", g.shape) * b2 + (g * g) * (f32(1) - b2)  # weighted_add, utils_mtf.py:332-333
        p1 = slot("exp_avg_p1", g.shape) * b1 + g * (f32(1) - b1)
        st["exp_avg_p2"], st["exp_avg_p1"] = keep(p2), keep(p1)
        return _opt_rsqrt(p2 * debias_momentum(b2)) * p1 * debias_momentum(b1)

    if opt == "adam":
        return adam()
    if opt == "novograd":  # optimizers.py:46-57
        if g.ndim == 0:
            return adam()
        p1_old, p2_old = slot("exp_avg_p1", g.shape), slot("exp_avg_p2", ())
        p1 = b1 * p1_old + g * _opt_rsqrt(p2_old)
        p2 = p2_old * b2 + _sum_f32(g * g) * (f32(1) - b2)
        st["exp_avg_p1"], st["exp_avg_p2"] = keep(p1), keep(np.asarray(p2, f32))
        return b1 * p1 + g * _opt_rsqrt(p2 * debias_momentum(b2))
    if opt == "sm3":  # optimizers.py:60-76
        nd = g.ndim
        if nd == 0:
            return adam()
        acc = slot("dim0", g.shape[:1]).reshape([-1] + [1] * (nd - 1))
        for i in range(1, nd):
            shape = [1] * nd
            shape[i] = -1
            acc = np.minimum(acc, slot(f"dim{i}", (g.shape[i],)).reshape(shape))
        nu = acc + g * g
        for i in range(nd):
            red = tuple(j for j in range(nd) if j != i)
            st[f"dim{i}"] = keep(nu.max(axis=red) if red else nu)
        return g * _opt_rsqrt(nu)
    if opt == "adaptive_clip":  # optimizers.py:79-84
        clip = f32(float(args[0]))
        with np.errstate(divide="ignore"):
            gn = np.minimum(f32(1) / np.sqrt(_sum_f32(g * g)), f32(1e6))
        wn = np.maximum(np.sqrt(_sum_f32(w * w)), f32(1e-3))
        return g * np.minimum(wn * gn * clip, f32(1))
    if opt == "l2norm_clip":  # optimizers.py:87-90
        clip = f32(float(args[0]))
        return g * clip * (f32(1) / np.sqrt(np.maximum(_sum_f32(g * g), clip ** f32(-2))))
    if opt == "global_l2norm_clip":  # optimizers.py:93-98: the sum runs over the RAW gradients of all variables
        clip = f32(float(args[0]))
        return g * clip * (f32(1) / np.sqrt(np.maximum(f32(hyper["global_sq"]), clip ** f32(-2))))
    if opt == "value_clip":  # optimizers.py:101-103
        clip = f32(float(args[0]))
        return np.maximum(np.minimum(g, clip), -clip)
    if opt == "gradient_centralisation":  # optimizers.py:106-107
        return g - f32(np.mean(g.astype(np.float64)))
    if opt == "weight_centralisation":  # optimizers.py:110-111
        return g + f32(np.mean(w.astype(np.float64)))
    if opt == "learning_rate":  # optimizers.py:114-115
        return g * f32(hyper["lr"])
    if opt == "momentum":  # optimizers.py:118-128
        mult, gmult, nesterov = f32(float(args[0])), f32(float(args[1])), bool(int(args[2]))
        new_state = mult * slot("momentum", g.shape) + g * gmult
        st["momentum"] = keep(new_state)
        return g + mult * new_state if nesterov else new_state
    if opt == "graft":  # optimizers.py:147-151: direction of g, magnitude of the inner optimizer's step
        inner = _apply_stage(args[0], list(args[1:]), g, w, st, hyper)
        with np.errstate(divide="ignore"):
            return g * (f32(1) / np.sqrt(_sum_f32(g * g))) * np.sqrt(_sum_f32(inner * inner))
    raise NotImplementedError(f"oracle: optimizer '{opt}'")


def optimizer_step(sh: Shapes, params, grads, opt_state, lr: float, step_count: int = 1, optimizer: str = None,
                   seen=None, assign: bool = True):
    """In-place update of params (fp32 slices) and opt_state, all arithmetic in fp32
    (optimizer_calculation_dtype).  `grads` already carry the dtype the graph would hand over.
    update(), src/optimizer/__init__.py:31-66: the stages of the optimizer string in order, the weight-decay rule,
    assign_sub.  step_count = global_step + 1 (`:96` with grad_accumulation 1); `optimizer` overrides sh.optimizer.

    Macro-batching (src/run/train.py:35-74): `seen` = the weights this macro-batch's graph read (ctx.value: the
    chained tensors of the previous macro-batch; default: the cast of the slices); assign=False is the "add" arm --
    optimizer slots still move on (assign_ctx's tensors replace the slot variables in params.variable_cache, so the
    next macro-batch reads them), the slices stay, and the function returns what the next macro-batch's graph reads
    in place of each variable: subtract(var.value, cast(update)) in the activation dtype (`__init__.py:65-66`)."""
    f32 = np.float32
    act_bf16 = sh.calculation_dtype == "bfloat16"
    st_bf16 = sh.optimizer_slice_dtype == "bfloat16"
    lr = f32(lr)
    chain = [o.split(":") for o in (optimizer or sh.optimizer).split("-")]
    hyper = {"lr": lr, "step_count": step_count, "beta1": getattr(sh, "opt_beta1", 0.9),
             "beta2": getattr(sh, "opt_beta2", 0.99),
             "global_sq": sum(float(np.sum(g.astype(np.float64) ** 2)) for g in grads.values()),
             "keep": (lambda v: round_bf16(np.asarray(v, f32))) if st_bf16 else (lambda v: np.asarray(v, f32))}
    act = round_bf16 if act_bf16 else (lambda v: np.asarray(v, f32))
    chained = {}
    for name, p in params.items():
        g = grads[name].astype(f32)
        w = act(p) if seen is None else np.asarray(seen[name], f32)  # var.value / the chained tensor
        st = opt_state[name]
        for opt, *args in chain:
            g = _apply_stage(opt, args, g, w, st, hyper)
        if "rezero" in name:  # __init__.py:46-47
            g = g * f32(getattr(sh, "rezero_lr_multiplier", 0.1))
        if is_large_tensor(name) and sh.weight_decay > 0:  # __init__.py:49-61
            g = g + w * lr * f32(sh.weight_decay)
        if assign:
            params[name] = (p - g).astype(f32)  # assign_sub on the fp32 slice, __init__.py:64
        else:
            chained[name] = act(act(p) - act(g))  # subtract(ctx.var, cast(grad, calculation_dtype)), __init__.py:66
    return None if assign else chained


def train_step_macro(sh: Shapes, params, opt_state, tok_x, tok_y, global_step: int, macro: int, dtype=np.float64):
    """get_train_model with macro_batching = macro (src/run/train.py:19-77): row b * macro + s belongs to slice s;
    every slice is differentiated against the weights chained from the previous slice's non-assign update; only the
    last slice's update reaches the slices.  Returns (mean loss, first loss, last loss, lr)."""
    act_bf16 = sh.calculation_dtype == "bfloat16"
    lr = learning_rate(sh, global_step)
    seen = None
    losses = []
    for s in range(macro):
        sx, sy = tok_x[s::macro], tok_y[s::macro]
        graph_params = seen if seen is not None else (
            {k: round_bf16(v) for k, v in params.items()} if act_bf16 else params)
        loss, acc, state = forward(sh, graph_params, sx, sy, dtype=dtype)
        grads = backward(sh, state)
        if act_bf16:
            grads = {k: round_bf16(v.astype(np.float32)) for k, v in grads.items()}
        losses.append(loss)
        seen = optimizer_step(sh, params, grads, opt_state, lr, step_count=global_step + 1, seen=seen,
                              assign=s == macro - 1)
    return float(np.mean(losses)), losses[0], losses[-1], lr


def train_step(sh: Shapes, params, opt_state, tok_x, tok_y, global_step: int, dtype=np.float64):
    """One full step: forward, backward, update.  Returns (loss, accuracy, grads, lr)."""
    loss, acc, state = forward(sh, params, tok_x, tok_y, dtype=dtype)
    grads = backward(sh, state)
    if sh.calculation_dtype == "bfloat16":
        grads = {k: round_bf16(v.astype(np.float32)) for k, v in grads.items()}
    lr = learning_rate(sh, global_step)
    optimizer_step(sh, params, grads, opt_state, lr, step_count=global_step + 1)
    return loss, acc, grads, lr


# ----------------------------------------------------------------------------------------------------------
# the other registered activations and the feed_forward layer
# ----------------------------------------------------------------------------------------------------------
def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def _softplus(x):
    return np.where(x > 20, x, np.log1p(np.exp(np.minimum(x, 20))))


# forward, and the backward THE REFERENCE uses: its custom operations for lecun_tanh / silu / mish / softsign
# (src/model/activation.py:13-155) -- silu's and mish's are not the analytic derivatives -- and autodiff for the rest
ACT_FWD = {
    "relu": lambda x: np.maximum(x, 0), "gelu": gelu_fwd,
    "sigmoid": _sigmoid, "tanh": np.tanh, "lecun_tanh": lambda x: np.tanh(x) + x * 0.1,
    "silu": lambda x: x * _sigmoid(x), "mish": lambda x: x * np.tanh(_softplus(x)),
    "mtf_mish": lambda x: x * np.tanh(_softplus(x)), "softsign": lambda x: x / (1 + np.abs(x)), "exp": np.exp,
}


def _mish_ref_bwd(x):
    gte = np.tanh(_softplus(x))
    return gte + 1.0 - gte * gte * x * _sigmoid(x)     # activation.py:41-44


ACT_BWD = {  # d/dx as the reference computes it, to be multiplied by dy
    "relu": lambda x: (x > 0).astype(x.dtype), "gelu": gelu_grad,
    "sigmoid": lambda x: _sigmoid(x) * (1 - _sigmoid(x)), "tanh": lambda x: 1 - np.tanh(x) ** 2,
    "lecun_tanh": lambda x: 1.1 - np.tanh(x) ** 2,            # activation.py:112-113
    "silu": lambda x: (x - 1) * _sigmoid(x) + 1,              # activation.py:77-79
    "mish": _mish_ref_bwd,
    "mtf_mish": lambda x: np.tanh(_softplus(x)) + x * (1 - np.tanh(_softplus(x)) ** 2) * _sigmoid(x),
    "softsign": lambda x: 1 / (1 + np.abs(x)) ** 2,           # activation.py:147-148
    "exp": np.exp,
}


def activated_linear_fwd(x, ws, act=None, glu=None):
    """activated_linear without norm (src/model/basic.py:47-55): x [T, K]; ws = [W1 (, W2 (, W3))], each [K, N];
    glu in (None, 'glu', 'glu_add').  Returns (out, cache)."""
    f = ACT_FWD[act] if act else (lambda v: v)
    z1 = x @ ws[0]
    out = f(z1)
    cache = {"x": x, "z1": z1}
    if glu:
        z2 = x @ ws[1]
        cache.update(z2=z2, a=out)
        out = out * _sigmoid(z2)
    if glu == "glu_add":
        z3 = x @ ws[2]
        cache["z3"] = z3
        out = out + f(z3)
    return out, cache


def activated_linear_bwd(dy, ws, cache, act=None, glu=None):
    """Returns (dx, [dW...]) with the reference's activation backward formulas."""
    d = ACT_BWD[act] if act else (lambda v: np.ones_like(v))
    x = cache["x"]
    grads = []
    dx = 0
    if glu:
        s = _sigmoid(cache["z2"])
        dz2 = dy * cache["a"] * s * (1 - s)
        da = dy * s
    else:
        da = dy
    dz1 = da * d(cache["z1"])
    grads.append(x.T @ dz1)
    dx = dx + dz1 @ ws[0].T
    if glu:
        grads.append(x.T @ dz2)
        dx = dx + dz2 @ ws[1].T
    if glu == "glu_add":
        dz3 = dy * d(cache["z3"])
        grads.append(x.T @ dz3)
        dx = dx + dz3 @ ws[2].T
    return dx, grads


def sample_tokens(sh: Shapes, params, token_x, initial_pos: int, temperature: float = 0.0,
                  end_iterations: typing.Optional[int] = None, uniform_fn=None, dtype=np.float64):
    """autoregressive_model, language branch (src/run/inference.py:76-134): every iteration runs the whole forward
    pass, adds log(-log(U)) * (-temperature) to the logits, takes the argmax over the vocabulary, shifts it one
    position to the right (zero fill) and writes the entry at `position` into the token window.
    uniform_fn(position) -> U of shape [batch, sequence, vocab] (the reference draws tf.random_uniform)."""
    tok = np.array(token_x, dtype=np.int64).reshape(token_x.shape[0], sh.sequence_length)
    end = sh.sequence_length if end_iterations is None else end_iterations
    pos = initial_pos
    while pos < end:
        trace = {}
        forward(sh, params, tok[:, :, None], tok[:, :, None], dtype=dtype, trace=trace)
        logits = trace["logits"].reshape(tok.shape[0], sh.sequence_length, -1).astype(np.float32)
        if temperature != 0:
            u = np.clip(uniform_fn(pos).astype(np.float32), np.float32(1e-9), np.float32(0.99999994))
            logits = logits + np.log(-np.log(u)) * np.float32(-temperature)
        out = logits.argmax(-1)
        shifted = np.concatenate([np.zeros_like(out[:, :1]), out[:, :-1]], 1)  # shift(+1, wrap=False)
        tok[:, pos] = shifted[:, pos]
        pos += 1
    return tok


def count_params(sh: Shapes) -> int:
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    per_depth = H * K * I + I * H * K2 + H * K2 * K + 2 * H * K2 + 6 * H * K
    return L * per_depth + 2 * H * S * S + V * I0 + I0 * F + F * V


def flops_per_token_fwd(sh: Shapes) -> float:
    """causal-aware algorithmic forward FLOPs per token."""
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    return L * (2 * (F * I + I * H * K2 + H * K2 * K) + 2 * 2 * F * (S + 1) / 2) + 2 * (I0 * F + F * V)
