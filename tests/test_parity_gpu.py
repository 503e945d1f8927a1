"""Model-level parity: the CUDA path (through the C-ABI) against oracle/ on identical seeded weights and tokens.

Tolerances are north_star's: fp32 <= 1e-5 relative on per-sub-block activations, every gradient and the one-step
loss; bf16 <= 2e-2.  Optimizer state layout / step counters are compared exactly where exactness is defined
(lr == 0 leaves every weight bit-unchanged)."""
import copy

import numpy as np
import pytest
import torch

from oracle import obst_oracle as orc
from tests.helpers import make_params, rel, to_np, variables_by_oracle_name

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()


def run_product_step(params, tok_x, tok_y, trace=False):
    from obst.run.train import get_train_model, read_scalars
    if trace:
        params.debug_trace = {}
    x = torch.from_numpy(tok_x).to(params.device)
    y = torch.from_numpy(tok_y).to(params.device)
    out = get_train_model(params, None, None, None, x, y, None, None, None, params.manual_global_step)
    scal = read_scalars(params, out[3])
    return out, scal


def test_tiny_fp32_full_step_parity(cuda):
    """BASELINE.json configs[0]: tiny mixer LM (2 layers, d_model 256, ctx 256, batch 8), fp32, one full step."""
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    sh = orc.Shapes(cfg)
    op = orc.init_params(sh, seed=0)
    tok_x, tok_y = orc.synthetic_tokens(sh)
    params = make_params(cfg, op)
    params.global_step = 2048  # lr = 0.01 * 2048/4096
    (_, _, lr, *_), scal = run_product_step(params, tok_x, tok_y, trace=True)
    # ---- oracle, float64 forward/backward (the error budget of the fp32 kernels is measured against it)
    trace = {}
    loss, acc, state = orc.forward(sh, op, tok_x, tok_y, dtype=np.float64, trace=trace)
    grads = orc.backward(sh, state, trace=trace)
    assert abs(scal["loss"] - loss) / loss < 1e-5
    assert abs(scal["accuracy"] - acc) < 1e-9
    assert abs(lr - orc.learning_rate(sh, 2048)) < 1e-12
    # per-sub-block activations (the two streams after each of the 2L sub-blocks) and their reconstructions
    dbg = params.debug_trace
    assert rel(to_np(dbg["src"]), trace["src"]) < 1e-5
    for i in range(sh.depth):
        for c in range(2):
            for k in (f"x1_{i}_{c}", f"x2_{i}_{c}", f"rx1_{i}_{c}", f"rx2_{i}_{c}"):
                assert rel(to_np(dbg[k]), trace[k]) < 1e-5, k
    assert rel(to_np(dbg["out"]), trace["out"]) < 1e-5
    # every gradient
    byname = variables_by_oracle_name(params, sh)
    for name, var in byname.items():
        assert rel(to_np(var.grad), grads[name]) < 1e-5, name
    assert float(torch.triu(byname["mix_a"].grad, diagonal=1).abs().max()) == 0.0
    # one optimizer step, fp32 oracle arithmetic on the oracle's own gradients
    ost = orc.init_opt_state(op)
    op2 = {k: v.copy() for k, v in op.items()}
    g32 = {k: v.astype(np.float32) for k, v in grads.items()}
    orc.optimizer_step(sh, op2, g32, ost, orc.learning_rate(sh, 2048))
    slots = params.optimizer_state.slots
    from obst.optimizer.backend import slot_name
    for name, var in byname.items():
        assert rel(to_np(var.data), op2[name]) < 1e-6, name
        upd, upd_ref = to_np(var.data) - op[name], op2[name] - op[name]
        assert rel(upd, upd_ref) < 1e-3, name  # the update is lr*(+-1.9) on step one: sign flips of ~0 gradients
        assert rel(to_np(slots[slot_name(params, var.name, "momentum")]), ost[name]["momentum"]) < 1e-3, name
        for i in range(len(var.dims)):
            assert rel(to_np(slots[slot_name(params, var.name, f"dim{i}")]), ost[name][f"dim{i}"]) < 1e-4, (name, i)
    assert params.global_step == 2049 and params.manual_global_step == 1


def test_lr_zero_first_step_leaves_weights_bit_identical(cuda):
    """global_step 0 with linear_warmup => lr == 0: weights must be bit-unchanged, state must change (quirk 4)."""
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=1, train_batch_size=2)
    sh = orc.Shapes(cfg)
    op = orc.init_params(sh, seed=1)
    tok_x, tok_y = orc.synthetic_tokens(sh)
    params = make_params(cfg, op)
    (_, _, lr, *_), _ = run_product_step(params, tok_x, tok_y)
    assert lr == 0.0
    byname = variables_by_oracle_name(params, sh)
    for name, var in byname.items():
        assert np.array_equal(to_np(var.data), op[name]), name
    from obst.optimizer.backend import slot_name
    m = params.optimizer_state.slots[slot_name(params, byname["win_0"].name, "momentum")]
    assert float(m.abs().max()) > 0


def test_three_steps_track_the_oracle_fp32(cuda):
    """Three consecutive steps (non-zero SM3 accumulators and momentum from step 2 on)."""
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=1, train_batch_size=4, sequence_length=64)
    sh = orc.Shapes(cfg)
    op = orc.init_params(sh, seed=2)
    tok_x, tok_y = orc.synthetic_tokens(sh)
    params = make_params(cfg, op)
    params.global_step = 1000
    ost = orc.init_opt_state(op)
    oparams = {k: v.copy() for k, v in op.items()}
    for step in range(3):
        (_, _, lr, *_), scal = run_product_step(params, tok_x, tok_y)
        loss, acc, grads, olr = orc.train_step(sh, oparams, ost, tok_x, tok_y, 1000 + step)
        assert abs(lr - olr) < 1e-12
        assert abs(scal["loss"] - loss) / loss < 2e-5, (step, scal["loss"], loss)
    byname = variables_by_oracle_name(params, sh)
    for name, var in byname.items():
        assert rel(to_np(var.data), oparams[name]) < 1e-5, name


def test_micro_batching_matches_single_pass(cuda):
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=1, train_batch_size=8, sequence_length=64)
    sh = orc.Shapes(cfg)
    op = orc.init_params(sh, seed=3)
    tok_x, tok_y = orc.synthetic_tokens(sh)
    results = []
    for mb in (None, 2):
        params = make_params(cfg, op, micro_batch_size=mb)
        params.global_step = 4096
        _, scal = run_product_step(params, tok_x, tok_y)
        byname = variables_by_oracle_name(params, sh)
        results.append((scal["loss"], {k: to_np(v.grad) for k, v in byname.items()}))
    assert abs(results[0][0] - results[1][0]) / results[0][0] < 1e-6
    for k in results[0][1]:
        assert rel(results[1][1][k], results[0][1][k]) < 1e-5, k


@pytest.mark.parametrize("backend", [1, 0])
def test_bf16_step_parity(cuda, backend):
    """bf16 activations / bf16 optimizer slices (the dtypes of configs/32*_mixer.json) on a reduced shape,
    checked against the float64 oracle: <= 2e-2.  backend 1 = SIMT GEMMs, 0 = tcgen05 GEMMs."""
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=2, train_batch_size=4, sequence_length=256, features_per_head=64, calculation_dtype="bfloat16",
               storage_dtype="bfloat16", optimizer_slice_dtype="bfloat16", slice_dtype="float32")
    sh = orc.Shapes(cfg)
    op = orc.init_params(sh, seed=4)
    # the graph reads bf16-rounded weights: give the oracle the same values
    op_seen = {k: orc.round_bf16(v) for k, v in op.items()}
    tok_x, tok_y = orc.synthetic_tokens(sh)
    params = make_params(cfg, op, gemm_backend=backend)
    params.global_step = 2048
    _, scal = run_product_step(params, tok_x, tok_y, trace=True)
    trace = {}
    loss, acc, state = orc.forward(sh, op_seen, tok_x, tok_y, dtype=np.float64, trace=trace)
    grads = orc.backward(sh, state, trace=trace)
    assert abs(scal["loss"] - loss) / loss < 2e-2
    dbg = params.debug_trace
    for i in range(sh.depth):
        for c in range(2):
            for k in (f"x1_{i}_{c}", f"x2_{i}_{c}", f"rx1_{i}_{c}", f"rx2_{i}_{c}"):
                assert rel(to_np(dbg[k]), trace[k]) < 2e-2, k
    byname = variables_by_oracle_name(params, sh)
    for name, var in byname.items():
        assert rel(to_np(var.grad), grads[name]) < 2e-2, name
