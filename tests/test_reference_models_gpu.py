"""The product against the REFERENCE'S OWN CODE, executed: tests/golden/reference_models.npz holds, for a zoo of toy
configs (every registered layer / memory strategy / macro-batching the product implements), what
src/run/train.py:get_train_model returned when /root/reference's Python ran on the NumPy stand-ins of oracle/ref_shim
(tests/golden/make_reference_vectors.py:run_model_case): the trainable variables in creation order with their full
names, the loss, the accuracy, the logits, every gradient, every updated variable.

Here the product builds the same config on the GPU (fp32, SIMT GEMMs), imports those weights IN CREATION ORDER, runs
its get_train_model and must reproduce all of it -- including the variable NAMES, which are the reference's scope
names.  Tolerances: both sides compute in float32 with different summation orders; the reversible strategies
differentiate at reconstructed activations, which amplifies that rounding on the early layers."""
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_models.npz")

CASES = ["mixer_depth2", "strategy_none", "strategy_checkpoint_skip", "strategy_none_skip", "feed_forward_glu",
         "activations", "cumsum_cummean", "group_linear", "transpose_sequence_features", "attention_embedded",
         "attention_context_biased_softmax", "attention_maps", "attention_shared_key_value",
         "attention_shared_key_value_maps", "attention_relative", "attention_relative_learned", "attention_axial",
         "attention_stack_revnet", "split_path", "initial_position_embedding", "macro_batching", "adam_chain"]


def _nested(npz, prefix):
    out = {}
    for k in npz.files:
        if k.startswith(prefix + "|"):
            node = out
            parts = k[len(prefix) + 1:].split("|")
            for p in parts[:-1]:
                node = node.setdefault(p, {})
            node[parts[-1]] = npz[k]
    return out


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()


@pytest.mark.parametrize("case", CASES)
def test_product_reproduces_the_reference_run(cuda, case):
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.run.train import get_train_model, read_scalars
    with np.load(GOLDEN) as f:
        rec = _nested(f, case)
    cfg = json.loads(str(rec["config"]))
    created = [str(n) for n in rec["created"]]
    params = ModelParameter(gpt_mode(dict(cfg, gemm_backend=1)))
    params.import_variables = [rec["init"][n] for n in created]
    params.global_step = 1500
    macro = int(cfg.get("macro_batching", 1))
    x = torch.from_numpy(rec["tok_x"]).to(params.device)
    y = torch.from_numpy(rec["tok_y"]).to(params.device)
    report = []
    prev = {n: rec["init"][n] for n in created}
    # gradients: first-hand under 'none' / 'checkpoint'; the reversible strategy differentiates at activations
    # reconstructed in fp32 (x1 = y1 - f(x2), src/model/revnet.py:89), which costs the first blocks ~1e-3.
    # updates: relative to the size of the step; sm3's g / sqrt(g^2) magnifies gradient noise on small entries
    reversible = cfg.get("memory_reduction_strategy", "revnet") == "revnet"
    gtol, tol = (3e-3, 2e-2) if reversible else (2e-4, 5e-3)
    for s in sorted(rec["steps"], key=int):
        step = rec["steps"][s]
        out = get_train_model(params, None, None, None, x, y, None, None, None, params.manual_global_step)
        scal = read_scalars(params, out[3])
        acc = read_scalars(params, out[8])["accuracy"] if out[8] is not None else scal["accuracy"]
        names = list(params.variable_cache.keys())
        assert names == created, [(a, b) for a, b in zip(names, created) if a != b][:3]
        loss = float(step["loss"])
        assert abs(scal["loss"] - loss) < 2e-5 * abs(loss), (case, s, scal["loss"], loss)
        assert abs(acc - float(step["accuracy"])) < 1e-6
        from obst.optimizer.optimizers import FusedPlan
        fused = isinstance(params.optimizer_state.plan, FusedPlan)
        errs = {}
        gmax = max(float(np.linalg.norm(g)) for g in step["grads"].values())
        for n in created:
            var = params.variable_cache[n]
            want_w = step["state"][n]
            if float(np.linalg.norm(step["grads"][n])) < 1e-6 * gmax:
                # the loss does not depend on this variable (the query projection of a `shared_key_value` attention,
                # whose output is rowsum(map) * key): its reference "gradient" is rounding noise, which the adaptive
                # optimizer turns into a step of arbitrary sign.  The product's gradient must be negligible too.
                assert float(var.grad.float().norm()) < 1e-5 * gmax, (case, n)
                prev[n] = want_w
                continue
            # (with macro-batching the product's buffer holds the last slice's gradient only; the generic optimizer
            # chains -- adam here, and the adam that rank-0 variables get under sm3 -- work in place on the buffer)
            if macro == 1 and fused and want_w.ndim > 0:
                eg = rel(var.grad.detach().float().cpu().numpy().reshape(step["grads"][n].shape), step["grads"][n])
                assert eg < gtol, (case, s, n, "gradient", eg)
            moved = float(np.linalg.norm(want_w.astype(np.float64) - prev[n].astype(np.float64)))
            got = var.data.detach().float().cpu().numpy().reshape(want_w.shape)
            # error of the UPDATE (what the step did to the variable), relative to the size of that update
            errs[n] = float(np.linalg.norm(got.astype(np.float64) - want_w.astype(np.float64))) / (moved + 1e-30)
            prev[n] = want_w
        worst = max(errs, key=errs.get)
        report.append((s, worst, errs[worst]))
        assert errs[worst] < tol, (case, s, worst, errs[worst])
    print(case, report)
