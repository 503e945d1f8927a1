"""The registered layers beside the mixer hot path, each through the layer registry exactly as a
block_config string would reach it, against a float64 statement of the reference's lines written with torch autograd
(these are small floating-point compositions; the statement cites the lines it follows).  fp32 at 1e-5."""
import copy

import numpy as np
import pytest
import torch

from oracle import obst_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def run_layers(layers, seq=16, fph=16, batch=2, seed=20, **cfg_overrides):
    """Runs one block made of `layers` on a random [batch, seq, heads, fph] input; returns everything a checker needs."""
    from obst import utils_core
    from obst.dataclass import BlockConfig, ModelParameter, gpt_mode
    from obst.model.backend import run_tape
    from obst.model.frontend import block_part_fn
    from obst.utils_core import scoped
    from obst.utils_mtf import Tensor
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=1, train_batch_size=batch, sequence_length=seq, features_per_head=fph)
    cfg.update(cfg_overrides)
    params = ModelParameter(gpt_mode(cfg))
    rng = np.random.default_rng(seed)
    H = params.heads
    x = torch.from_numpy(rng.standard_normal((batch, seq, H, fph)).astype(np.float32)).cuda()
    dims = [params.batch_dim._replace(size=batch), params.sequence_dim] + params.feature_dims
    utils_core.NAME_INDICES.clear()
    params.cached_parameters.clear()
    params.tape = []
    block = BlockConfig({"layer": layers}, "revnet")
    out = scoped("gpt", lambda: scoped("body", lambda: block_part_fn(params, block, Tensor(x, dims), "0")))
    tape, params.tape = params.tape, None
    dy = torch.from_numpy(rng.standard_normal(tuple(out.data.shape)).astype(np.float32)).cuda()
    dx = run_tape(tape, dy)
    return params, x, out, dy, dx, list(params.variable_cache.values())


def test_group_linear_and_reduced_half_linear(cuda):
    """group_linear (src/model/basic.py:72-74): out[b,s,h,j] = sum_k x[b,s,h,k] W[h,k,j];
    reduced_half_linear (:89-90): the same weights applied to the head SUM, out[b,s,h,j] = sum_k (sum_h' x)[k] W[h,k,j]."""
    params, x, out, dy, dx, variables = run_layers(["group_linear"])
    (w,) = variables
    xr = x.double().requires_grad_(True)
    wr = w.data.double().requires_grad_(True)
    ref = torch.einsum("bshk,hkj->bshj", xr, wr)
    ref.backward(dy.double())
    assert rel(out.data, ref.detach()) < 1e-5 and rel(dx, xr.grad) < 1e-5 and rel(w.grad, wr.grad) < 1e-5
    params, x, out, dy, dx, variables = run_layers(["reduced_half_linear"])
    (w,) = variables
    xr = x.double().requires_grad_(True)
    wr = w.data.double().requires_grad_(True)
    ref = torch.einsum("bsk,hkj->bshj", xr.sum(2), wr)
    ref.backward(dy.double())
    assert tuple(out.data.shape) == tuple(ref.shape)
    assert rel(out.data, ref.detach()) < 1e-5 and rel(dx, xr.grad) < 1e-5 and rel(w.grad, wr.grad) < 1e-5


def test_sum_heads_and_transpose_sequence_features(cuda):
    """sum_heads (src/model/basic.py:77-78) = reduce_sum over heads; transpose_sequence_features (:81-86) exchanges
    the sequence and features_per_head axes (equal sizes)."""
    params, x, out, dy, dx, _ = run_layers(["sum_heads"])
    assert rel(out.data, x.double().sum(2)) < 1e-6
    assert rel(dx, dy.double()[:, :, None, :].expand_as(x)) == 0.0
    params, x, out, dy, dx, _ = run_layers(["transpose_sequence_features"], seq=16, fph=16)
    assert torch.equal(out.data, x.permute(0, 3, 2, 1).contiguous())
    assert torch.equal(dx, dy.permute(0, 3, 2, 1).contiguous())


def test_rezero_layer_and_its_optimizer_rule(cuda):
    """rezero (src/model/basic.py:21-22): x * s, s a rank-0 variable initialised to 0; ds = sum(x * dy).  The
    optimizer gives rank-0 variables adam in place of sm3 (src/optimizer/optimizers.py:61-62) and rezero variables
    the extra learning-rate multiplier (src/optimizer/__init__.py:46-47)."""
    params, x, out, dy, dx, variables = run_layers(["group_linear", "rezero"])
    w, s = variables
    assert s.data.dim() == 0 and float(s.data) == 0.0
    assert float(out.data.abs().max()) == 0.0 and float(dx.abs().max()) == 0.0
    h = torch.einsum("bshk,hkj->bshj", x.double(), w.data.double())
    assert abs(float(s.grad) - float((h * dy.double()).sum())) < 1e-5 * float((h * dy.double()).abs().sum())
    # one optimizer step on this single scalar through the product's optimizer
    from obst.optimizer import _ensure_state
    state = _ensure_state(params)
    assert state.scalar_opt is not None and len(state.variables) == 1
    g = float(s.grad)
    lr = 0.01
    state.step(lr, refresh_transposed=False, step_count=1)
    torch.cuda.synchronize()
    # oracle: adaptive_clip -> adam (rank 0) -> momentum:0.9:1:1 -> learning_rate, then the rezero multiplier
    st = {}
    f32 = np.float32
    hyper = {"lr": f32(lr), "step_count": 1, "beta1": 0.9, "beta2": 0.99, "global_sq": g * g,
             "keep": lambda v: np.asarray(v, f32)}
    gg = np.asarray(g, f32)
    w = np.asarray(0.0, f32)
    for opt, *a in (o.split(":") for o in orc.TINY_CONFIG["optimizer"].split("-")):
        gg = orc._apply_stage(opt, a, gg, w, st, hyper)
    expect = -float(gg) * params.rezero_lr_multiplier
    assert abs(float(s.data) - expect) < 1e-5 * abs(expect), (float(s.data), expect)


def test_split_path_add_and_multiply(cuda):
    """split_path (src/model/frontend.py:39-55): every ';' path is applied to the same input, outputs summed or
    multiplied."""
    for mode in ("add", "multiply"):
        params, x, out, dy, dx, variables = run_layers([f"split_path-{mode};cumsum;group_linear,activation-tanh;cummean"])
        (w,) = variables
        xr = x.double().requires_grad_(True)
        wr = w.data.double().requires_grad_(True)
        p1 = torch.cumsum(xr, 1)
        p2 = torch.tanh(torch.einsum("bshk,hkj->bshj", xr, wr))
        p3 = torch.cumsum(xr, 1) / (1 + torch.arange(x.shape[1], dtype=torch.float64, device=xr.device)).view(1, -1, 1, 1)
        ref = p1 + p2 + p3 if mode == "add" else p1 * p2 * p3
        ref.backward(dy.double())
        assert rel(out.data, ref.detach()) < 1e-5, mode
        assert rel(dx, xr.grad) < 1e-5 and rel(w.grad, wr.grad) < 1e-5, mode


def test_dropout_keeps_the_stated_fraction_and_replays_its_mask(cuda):
    """dropout-dropout_rate0.25 (src/model/basic.py:25-30 -> mtf.dropout): inverted dropout; the reference's own test
    (tests/basic_pointwise_test.py) checks the kept fraction, so does this, plus: survivors are scaled by 1/keep and
    the backward pass applies the same mask."""
    params, x, out, dy, dx, _ = run_layers(["dropout-dropout_rate0.25"], seq=64, fph=64, batch=4)
    kept = out.data != 0
    frac = float(kept.float().mean())
    assert abs(frac - 0.75) < 0.01, frac
    assert torch.allclose(out.data[kept], x[kept] / 0.75, rtol=1e-6)
    assert torch.equal(dx != 0, kept & (dy != 0)) and torch.allclose(dx[kept], dy[kept] / 0.75, rtol=1e-6)
    # evaluation mode (params.train False) and rate 0: identity
    params, x, out, dy, dx, _ = run_layers(["dropout-dropout_rate0.25"], train=False)
    assert torch.equal(out.data, x)


def _attention_reference(layer, x, variables, params, dy, pos_builder=None):
    """float64 torch statement of src/model/spatial.py:42-81 for the flag set in `layer`; returns (out, dx, grads)."""
    flags = layer.split('-')[1:]
    B, S, H, K = x.shape
    it = iter(variables)
    leaves = []

    def nxt():
        v = next(it)
        t = (v.data if hasattr(v, "full_name") else v.value).double().clone().requires_grad_(hasattr(v, "full_name"))
        leaves.append((v, t))
        return t

    xr = x.double().requires_grad_(True)
    m = torch.tril(torch.ones(S, S, dtype=torch.float64, device=x.device))
    masked = 0 in params.masked_attention_dimensions
    mm = m if masked else torch.ones_like(m)
    base = None
    if "dot_product" in flags or "input_as_value" not in flags:
        base = torch.einsum("bshk,hki->bsi", xr, nxt())
        if "in:relu" in flags:
            base = base.relu()
    logit = None
    val = None
    if "dot_product" in flags:
        key = 0
        if "embedded" in flags or "context" in flags:
            key = torch.einsum("bsi,ihk->bshk", base, nxt())
        if "embedded" in flags or "positional" in flags:
            key = key + (pos_builder(nxt) if pos_builder is not None else nxt())[None]
        qry = torch.einsum("bsi,ihk->bshk", base, nxt()) * S ** -0.5
        if key.dim() == 3:
            key = key[None].expand(B, -1, -1, -1)
        logit = torch.einsum("bshk,bthk->bsht", qry, key)
        if "shared_key_value" in flags:
            val = key
    if "biased_softmax" in flags:
        bias = nxt() * mm[None]
        logit = bias.permute(1, 0, 2)[None] + (logit if logit is not None else 0)
    logit = logit.masked_fill(m[None, :, None, :] == 0, float("-inf"))
    p = torch.softmax(logit, -1)
    if "biased_attention_map" in flags:
        p = p + (nxt() * mm[None]).permute(1, 0, 2)[None]
    if "scale_attention_map" in flags:
        p = p * (nxt() * mm[None]).permute(1, 0, 2)[None]
    if val is None:
        val = xr if "input_as_value" in flags else torch.einsum("bsi,ihk->bshk", base, nxt())
    if "dot_product" in flags and "shared_key_value" in flags:
        # src/model/spatial.py:63-64,76-81: `val = key` skips the rename of its sequence dim, so the last einsum
        # contracts only the map's `_sequence`: out = rowsum(map) * key -- no token mixing.  Confirmed by running the
        # reference's own code (tests/golden/reference_models.npz, case attention_shared_key_value).
        out = p.sum(-1)[..., None] * val
    else:
        out = torch.einsum("bsht,bthk->bshk", p, val)
    out.backward(dy.double())
    return out.detach(), xr.grad, [(v, t.grad) for v, t in leaves if t.requires_grad]


@pytest.mark.parametrize("layer", [
    "attention-dot_product-embedded-absolute-in:relu",
    "attention-dot_product-context-input_as_value-biased_softmax-absolute",
    "attention-biased_softmax-scale_attention_map-biased_attention_map-input_as_value-absolute",
    "attention-dot_product-positional-relative-input_as_value",
    "attention-dot_product-embedded-shared_key_value-absolute",
])
def test_dot_product_attention_variants(cuda, layer):
    """The general attention layer (src/model/spatial.py:42-81): dot-product logits from projected queries / keys
    (+ absolute or sinusoidal relative position keys), `biased_softmax`, the causal soft-max, `biased_attention_map`,
    `scale_attention_map`, projected / shared / input values -- output, input gradient and every weight gradient
    against a float64 statement of those lines, fp32 at 1e-5."""
    # (the sinusoidal embedding takes exp() of the feature index, src/model/embedding.py:167-169: more than ~88
    # features overflow fp32 in the reference as well, so that variant runs with 8 heads x 8 features)
    fph = 8 if "relative" in layer.split('-') else 16
    params, x, out, dy, dx, variables = run_layers([layer], seq=24, fph=fph, batch=2, position_embedding="relative")
    # constants (relative embeddings) are not in the variable cache: rebuild the creation-order list
    from obst.model.embedding import Constant, relative_embedding  # noqa: F401
    ordered = list(variables)
    if "relative" in layer.split('-'):
        from obst.dataclass import BlockArgs
        from obst.utils_mtf import Tensor as T_
        dims = [params.batch_dim._replace(size=2), params.sequence_dim] + params.feature_dims
        const = Constant(relative_embedding(BlockArgs(params, T_(x, dims), layer.split('-')[1:]),
                                            [params.sequence_dim] + params.feature_dims),
                         [params.sequence_dim] + params.feature_dims)
        ordered.insert(1, const)  # created after the input projection, before the query projection
    ref, dx_ref, grads = _attention_reference(layer, x, ordered, params, dy)
    assert rel(out.data, ref) < 1e-5
    assert rel(dx, dx_ref) < 1e-5
    for v, g in grads:
        if g is None or float(g.abs().max()) < 1e-12:  # the output does not depend on it (query projection under shared_key_value)
            assert float(v.grad.abs().max()) < 1e-6, v.name  # soft-max backward of a row-constant gradient: rounding noise
            continue
        assert rel(v.grad, g.reshape(v.grad.shape)) < 1e-5, v.name


def _sinusoid(params, x, flags):
    """The sinusoidal relative embedding (src/model/embedding.py:128-172) as the product builds it, in float64.  It is
    taken from the product: the schedule is exp(feature index), whose sine is rounding noise in fp32 beyond the first
    few features -- in the reference as here -- so a float64 restatement would not be the same constant."""
    from obst.dataclass import BlockArgs
    from obst.model.embedding import relative_embedding
    from obst.utils_mtf import Tensor as T_
    dims = [params.batch_dim._replace(size=x.shape[0]), params.sequence_dim] + params.feature_dims
    return relative_embedding(BlockArgs(params, T_(x, dims), flags), [params.sequence_dim] + params.feature_dims).double()


@pytest.mark.parametrize("kind", ["axial", "axial-3", "relative-learned"])
def test_composed_position_embeddings_in_attention(cuda, kind):
    """`axial` (the outer product of per-factor embeddings, src/model/embedding.py:192-212) and `relative-learned` (the
    sinusoid times a learned feature scale, :214-217) as the position keys of dot-product attention: output, input
    gradient and the gradients of the factor / scale variables against float64 autograd, fp32 at 1e-5."""
    layer = f"attention-dot_product-positional-{kind}-input_as_value"
    S, fph = 24, 8 if "relative" in kind else 16
    params, x, out, dy, dx, variables = run_layers([layer], seq=S, fph=fph, batch=2)
    H = params.heads

    def pos_builder(nxt):
        if kind == "relative-learned":
            return _sinusoid(params, x, layer.split('-')[1:]) * nxt()[None]
        splits = 3 if kind.endswith("3") else 2
        parts = [nxt() for _ in range(splits)]
        sizes = [p.shape[0] for p in parts]
        assert int(np.prod(sizes)) == S, sizes
        full = None
        for i, p in enumerate(parts):
            v = p.view([sizes[j] if j == i else 1 for j in range(splits)] + [H, fph])
            full = v if full is None else full * v
        return full.reshape(S, H, fph)

    ref, dx_ref, grads = _attention_reference(layer, x, list(variables), params, dy, pos_builder)
    assert rel(out.data, ref) < 1e-5
    assert rel(dx, dx_ref) < 1e-5
    assert len(grads) == len(variables)
    for v, g in grads:
        assert rel(v.grad, g.reshape(v.grad.shape)) < 1e-5, v.name


@pytest.mark.parametrize("kind", ["absolute", "axial", "relative", "relative-learned"])
def test_initial_position_embedding(cuda, kind):
    """use_initial_position_embedding (src/model/__init__.py:97-99): src + embed(position_embedding, [sequence] +
    feature_dims) in front of the body, with the gradient reaching the embedding's variables."""
    from obst import utils_core
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.model import _initial_position_embedding
    from obst.model.backend import run_tape
    from obst.utils_core import scoped
    from obst.utils_mtf import Tensor
    S, fph, B = 24, 8, 3
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=1, train_batch_size=B, sequence_length=S, features_per_head=fph, position_embedding=kind,
               use_initial_position_embedding=True)
    params = ModelParameter(gpt_mode(cfg))
    H = params.heads
    rng = np.random.default_rng(3)
    x = torch.from_numpy(rng.standard_normal((B, S, H, fph)).astype(np.float32)).cuda()
    dy = torch.from_numpy(rng.standard_normal((B, S, H, fph)).astype(np.float32)).cuda()
    dims = [params.batch_dim._replace(size=B), params.sequence_dim] + params.feature_dims
    utils_core.NAME_INDICES.clear()
    params.cached_parameters.clear()
    params.tape = []
    out = scoped("gpt", lambda: _initial_position_embedding(params, Tensor(x, dims)))
    tape, params.tape = params.tape, None
    dx = run_tape(tape, dy)
    variables = list(params.variable_cache.values())
    leaves = [v.data.double().clone().requires_grad_(True) for v in variables]
    if kind == "absolute":
        pos = leaves[0]
    elif kind == "relative":
        pos = _sinusoid(params, x, kind.split('-'))
    elif kind == "relative-learned":
        pos = _sinusoid(params, x, kind.split('-')) * leaves[0][None]
    else:
        a, b = leaves
        pos = (a[:, None] * b[None]).reshape(S, H, fph)
    ref = x.double() + pos[None]
    assert rel(out.data, ref.detach()) < 1e-6
    assert torch.equal(dx, dy)
    if leaves:
        ref.backward(dy.double())
        for v, leaf in zip(variables, leaves):
            assert rel(v.grad, leaf.grad.reshape(v.grad.shape)) < 1e-5, v.name
    else:
        assert not variables
