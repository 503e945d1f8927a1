"""Host-side mirror of the reference interface: config surface, layer/optimizer/LR registries, named-dimension
shape rules, scope naming, weight-decay predicate.  CPU only (no kernel is launched)."""
import copy

import numpy as np
import pytest

from oracle import obst_oracle as orc
from obst.configs import mixer_config
from obst.dataclass import BlockArgs, ModelParameter, gpt_mode
from obst.utils_mtf import Dimension, Tensor, anonymize_dim, linear_shapes


class FakeData:
    """shape-only stand-in for a device buffer (the shape algebra never touches memory)."""

    def __init__(self, shape):
        self.shape = tuple(shape)

    def numel(self):
        return int(np.prod(self.shape))


def fake(dims):
    return Tensor(FakeData([d.size for d in dims]), dims)


def params_for(name="32big_mixer", **kw):
    return ModelParameter(mixer_config(name, **kw))


def test_reference_json_surface(capsys):
    cfg = mixer_config("32big_mixer")
    cfg.update(adaptive_gradient_clipping=False, gradient_clip=1)  # present in configs/*.json, unknown to the code
    p = ModelParameter(gpt_mode(cfg))
    out = capsys.readouterr().out
    assert "WARNING: Unknown ModelParameter adaptive_gradient_clipping=False" in out
    assert "WARNING: Unknown ModelParameter gradient_clip=1" in out
    assert p.features == 4096 and p.intermediate[0].size == 512
    assert p.mesh_shape == "b:4,h:8" and p.layout == "batch:b,heads:h"
    assert str(p.calculation_dtype) == "torch.bfloat16" and str(p.slice_dtype) == "torch.float32"
    assert p.use_video is False
    assert p.learning_rate_config["linear_warmup"].final_step == 4096
    assert [c.memory_reduction_strategy for c in p.block_config] == ["revnet", "revnet"]


@pytest.mark.parametrize("bad,err", [
    ({"grad_accumulation": 2}, "Gradient accumulation"),
    ({"use_language": False, "use_video": False}, "No model can be built"),
    ({"features_per_head": None, "features": None}, "Either features or features_per_head"),
    ({"calculation_dtype": "float16"}, "supported dtypes"),
])
def test_config_errors_match_the_reference(bad, err):
    cfg = mixer_config("tiny")
    cfg.update(bad)
    with pytest.raises(ValueError, match=err):
        ModelParameter(cfg)


def test_blockargs_quirks():
    p = params_for("tiny")
    a = BlockArgs(p, None, ["x", "y"], is_last=True)
    b = a(["z"])
    assert b.name_extras == ["z"] and b.is_last is False  # is_last is dropped: scale_by_depth never fires
    c = a("w")
    assert c.name_extras[-1] is str  # the reference appends the TYPE, not the string
    assert "x" in a and len(a) == 2 and a[1] == "y"
    with pytest.raises(ValueError):
        a(3.5)


@pytest.mark.parametrize("name", ["32big_mixer", "32ctx_mixer", "32mixer_group", "tiny"])
def test_linear_shapes_give_the_survey_weight_shapes(name):
    p = params_for(name)
    sh = orc.Shapes(mixer_config(name))
    B, S, H, K, F, I, K2, I0, V, L = sh.dims
    b, s = Dimension("batch", 2), p.sequence_dim
    x = fake([b, s] + p.feature_dims)
    # in: (extras ['relu'])   [heads, key] -> [intermediate]
    old, new = linear_shapes(BlockArgs(p, x, ["relu"]))
    assert [d.size for d in old + new] == [H, K, I]
    # mid: (extras [..., 'group']) tensor [b, s, intermediate]
    h1 = fake([b, s] + p.intermediate)
    old, new = linear_shapes(BlockArgs(p, h1, ["relu", "norm", "shift", "scale", "features", "group"]))
    assert [d.size for d in old + new] == [I, H, K2]
    assert new[1] == anonymize_dim(p.key_dim, K2)
    # out: (extras ['group']) tensor [b, s, heads, _key(K2)] : heads is in old AND new -> grouped
    h2 = fake([b, s, p.head_dim, anonymize_dim(p.key_dim, K2)])
    old, new = linear_shapes(BlockArgs(p, h2, ["group"]))
    assert [d.size for d in old] == [H, K2] and [d.size for d in new] == [H, K]
    # norm widths
    assert [d.size for d in linear_shapes(BlockArgs(p, x, ["shift", "scale", "features", "group"])).old] == [H, K]
    assert [d.size for d in linear_shapes(BlockArgs(p, h2, ["norm", "group"])).old] == [H, K2]


def test_registries_carry_the_reference_names():
    from obst.model.frontend import LAYER_FUNCTIONS
    from obst.optimizer.learning_rate import MODULES
    from obst.optimizer.optimizers import OPTIMIZERS
    assert set(LAYER_FUNCTIONS) == {'feed_forward', 'attention', 'cummean', 'cumsum', 'norm', 'rezero', 'activation',
                                    'convolution', 'dropout', 'group_linear', 'split_path',
                                    'feed_forward_product_key_memory', 'product_key_memory', 'reduced_half_linear',
                                    'transpose_sequence_features', 'bottleneck_group_linear', 'sum_heads'}
    assert set(OPTIMIZERS) == {"adam", "sm3", "novograd", "adaptive_clip", "l2norm_clip", "value_clip",
                               "gradient_centralisation", "weight_centralisation", "learning_rate",
                               "global_l2norm_clip", "momentum", "graft"}
    assert set(MODULES) == {"linear_warmup", "exponential_decay", "linear_decay", "lower_bound", "upper_bound"}
    with pytest.raises(NotImplementedError):
        LAYER_FUNCTIONS["convolution"](None)  # registered by the reference, outside the scope of this path


def test_optimizer_string_lowers_to_the_fused_plan():
    from obst.optimizer.context import OptimizerCtx
    from obst.optimizer.optimizers import OPTIMIZERS, lower_plan
    p = params_for("tiny")

    def plan_of(s):
        ctx = OptimizerCtx(p, [], 0.0, 0, True)
        for opt in s.split('-'):
            opt, *args = opt.split(':')
            OPTIMIZERS[opt](ctx, *args)
        return lower_plan(ctx.plan)

    f = plan_of("adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate")
    assert (f.clip, f.beta, f.nesterov) == (0.003, 0.9, True)
    f = plan_of("sm3-momentum:0.5:1:0-learning_rate")
    assert (f.clip, f.beta, f.nesterov) == (0.0, 0.5, False)
    # every other chain of registered stages lowers to the stage-by-stage plan, arguments checked up front
    from obst.optimizer.optimizers import GenericPlan
    g = plan_of("learning_rate")
    assert isinstance(g, GenericPlan) and g.stages == (("learning_rate", ()),)
    g = plan_of("adam-learning_rate")
    assert g.stages == (("adam", ()), ("learning_rate", ()))
    g = plan_of("sm3-momentum:0.9:0.5:1-learning_rate")   # gradient multiplier != 1: not the fused kernel's form
    assert isinstance(g, GenericPlan) and g.stages[1] == ("momentum", ("0.9", "0.5", "1"))
    g = plan_of("global_l2norm_clip:1.0-graft:adam-momentum:0.9:1:0-learning_rate")
    assert g.stages[1] == ("graft", ("adam",))
    with pytest.raises(TypeError):
        plan_of("momentum:0.9-learning_rate")
    with pytest.raises(ValueError):
        plan_of("graft:nothing")
    with pytest.raises(KeyError):
        plan_of("sgd")


@pytest.mark.parametrize("step", [0, 1, 100, 2048, 4095, 4096, 100000])
def test_learning_rate_modules_match_the_oracle(step):
    from obst.optimizer.learning_rate import get_learning_rate
    cfg = mixer_config("tiny")
    cfg["learning_rate_config"] = {"linear_warmup": {"final_step": 4096},
                                   "exponential_decay": {"start_step": 5000, "factor": 0.9999},
                                   "lower_bound": {"factor": 1e-6}}
    p = ModelParameter(cfg)
    p.global_step = step
    sh = orc.Shapes(cfg)
    assert np.isclose(float(get_learning_rate(p).learning_rate), orc.learning_rate(sh, step), rtol=1e-6, atol=0)


def test_scope_names_and_snapshot_restore():
    from obst import utils_core
    utils_core.NAME_INDICES.clear()
    names = []
    utils_core.scoped("body", lambda: utils_core.scoped("0_", lambda: names.append(utils_core.current_scope())))
    snap = utils_core.snapshot()
    utils_core.scoped("0_", lambda: names.append(utils_core.current_scope()))
    utils_core.restore(snap)
    utils_core.scoped("0_", lambda: names.append(utils_core.current_scope()))
    assert names == ["body0v/0_0v", "0_1v", "0_1v"]
    assert utils_core.current_scope() == ""


def test_weight_decay_predicate_on_reference_style_names():
    """src/optimizer/__init__.py:49-57: only the three feed-forward weights decay."""
    from obst.optimizer import is_large_tensor
    p = params_for("32big_mixer")
    H, K = p.head_dim, p.key_dim
    I = p.intermediate[0]
    K2 = anonymize_dim(K, K.size * 2)

    class V:
        def __init__(self, name, dims):
            self.name, self.dims = name, dims
            self.size = int(np.prod([d.size for d in dims]))

    body = "gpt0v/body0v/3_0v/bottleneck_group_linear_3v/orthogonal_var9v/get_variable9"
    assert is_large_tensor(p, V(body, [H, K, I]))
    assert is_large_tensor(p, V(body, [I, H, K2]))
    assert is_large_tensor(p, V(body, [H, K2, K]))
    assert not is_large_tensor(p, V("gpt0v/body0v/3_0v/norm_6v/normal_var12v/get_variable12", [H, K]))
    s, s_ = p.sequence_dim, anonymize_dim(p.sequence_dim)
    assert not is_large_tensor(p, V("gpt0v/body0v/0_1v/attention_0v/embed0v/normal_var4v/get_variable4", [H, s, s_]))
    assert not is_large_tensor(p, V("gpt0v/input0v/orthogonal_var0v/get_variable1", [Dimension("p", 1), I, H, K]))
    assert not is_large_tensor(p, V("gpt0v/output0v/embed2v/orthogonal_var96v/get_variable96", [H, K, p.vocab_dim]))


def test_shard_rows():
    from obst.parallel import shard_rows
    assert [shard_rows(1024, r, 8) for r in (0, 7)] == [(0, 128), (896, 1024)]
    with pytest.raises(ValueError):
        shard_rows(10, 0, 4)


def test_analysis_matches_oracle_counts():
    from obst.analysis import count_params, flops_per_token_fwd
    for name in ("32big_mixer", "32ctx_mixer", "32mixer_group", "tiny"):
        p = params_for(name)
        sh = orc.Shapes(mixer_config(name))
        assert count_params(p) == orc.count_params(sh)
        assert flops_per_token_fwd(p) == orc.flops_per_token_fwd(sh)


def test_update_work_list_covers_every_row_once():
    """The optimizer update's own work list (graded in size towards the end so that the resident blocks run dry
    together, obst/optimizer/__init__.py:_drain_list): whatever the grading, every row of every tensor is covered
    exactly once, in the order of the list it was cut from, with at least 8 rows (one per warp) per item."""
    import os
    from obst.optimizer import _drain_list
    cols_of = [512, 1024, 512, 8, 256]
    rows_of = [4096, 4096, 8192, 8, 333]
    items = []
    for ti, rows in enumerate(rows_of):
        rpi = max(1, 131072 // min(cols_of[ti], 512))
        items += [(ti, r0, min(rpi, rows - r0)) for r0 in range(0, rows, rpi)]
    old = os.environ.get("OBST_OPT_DRAIN")
    try:
        for setting in ("0:1:1", "0.05:3:6", "0.3:2:8", "1.0:4:16"):
            os.environ["OBST_OPT_DRAIN"] = setting
            out = _drain_list(items, cols_of)
            covered = {ti: [] for ti in range(len(rows_of))}
            for ti, r0, rows in out:
                assert rows >= 1
                covered[ti].append((r0, rows))
            for ti, rows in enumerate(rows_of):
                pos = 0
                for r0, n in covered[ti]:
                    assert r0 == pos, (setting, ti, r0, pos)
                    pos += n
                assert pos == rows
            assert [t for t, _, _ in out] == sorted(t for t, _, _ in out)   # order of the tensors is kept
            if setting == "0:1:1":
                assert out == items
            else:
                assert len(out) > len(items)
    finally:
        if old is None:
            os.environ.pop("OBST_OPT_DRAIN", None)
        else:
            os.environ["OBST_OPT_DRAIN"] = old


def test_the_loss_rides_on_the_lm_head_only_where_its_epilogue_exists():
    """`_fused_loss` (obst/model/__init__.py): obst_gemm epilogue 12 needs bf16, the tcgen05 kernel, a 256-wide
    vocabulary in one tile, more than one 128-row tile of tokens, and a step that wants the gradient -- anything
    else keeps LM head and obst_xent_fwd_bwd apart."""
    import torch

    from obst.model import _fused_loss
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(calculation_dtype="bfloat16", train_batch_size=2, sequence_length=128)

    def decide(dtype=torch.bfloat16, tokens=256, tape=True, **overrides):
        c = copy.deepcopy(cfg)
        c.update(overrides)
        params = ModelParameter(gpt_mode(c))
        params.tape = [] if tape else None
        feat = [d.size for d in params.feature_dims]
        out = Tensor(torch.zeros([tokens] + feat, dtype=dtype), [Dimension("tokens", tokens)] + params.feature_dims)
        tgt = Tensor(torch.zeros(tokens, 1, dtype=torch.int32), [Dimension("tokens", tokens), params.token_patch_dim])
        return _fused_loss(params, out, tgt)

    fused = decide()
    assert fused is not None and fused.n_global == 256 and fused.sums.dtype == torch.float64
    assert abs(fused.z_loss - cfg.get("z_loss", 1e-4)) < 1e-12
    assert decide(dtype=torch.float32) is None            # fp32 runs the SIMT GEMM
    assert decide(tokens=128) is None                     # one 128-row tile: the 1-CTA kernel has no such epilogue
    assert decide(tape=False) is None                     # evaluation wants the logits, not their gradient
    assert decide(vocab_size=128) is None and decide(vocab_size=512) is None
    assert decide(keep_logits=True) is None
    assert decide(gemm_backend=1) is None
    assert decide(fuse_lm_head_loss=False) is None
