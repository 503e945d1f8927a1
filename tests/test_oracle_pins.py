"""Pins oracle/obst_oracle.py against vectors recorded from the REFERENCE'S OWN PYTHON (tests/golden/reference_*.npz,
made by tests/golden/make_reference_vectors.py in the build container: /root/reference's src.* imported unmodified,
its two third-party tensor libraries replaced by the eager NumPy stand-ins of oracle/ref_shim).  What is pinned is
the reference's op order, constants and rules -- its code, executed -- not XLA's kernels.  CPU only; nothing here
reads /root/reference."""
import copy
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import obst_oracle as orc  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def small_config():
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(sequence_length=16, features_per_head=8, heads=4, train_batch_size=2, vocab_size=32)
    return cfg


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


# the recorded variables (make_reference_vectors.variable_specs), in order, under the oracle's names: the name decides
# the oracle's weight-decay predicate and its rezero rule exactly as the scope name does in the reference
ORACLE_NAMES = ["win_0", "wout_0", "g0_0", "mix_a", "w0", "rezero_0"]
FIRST_STEP = 1500

OPT_CASES = ["hot_path", "no_clip_heavyball", "adam", "novograd", "clips", "global_clip", "graft", "macro2",
             "hot_path_bf16", "adam_bf16"]


@pytest.mark.parametrize("case", OPT_CASES)
def test_optimizer_matches_the_reference_get_optimizer(case):
    """src/optimizer/__init__.py:get_optimizer (learning-rate context, step / beta / step_count arithmetic, every
    stage of src/optimizer/optimizers.py, the rezero multiplier, the weight-decay rule, assign_sub -- and for
    `macro2` the "add" / "update" call pair of src/run/train.py:35-74) against orc.optimizer_step on the same
    variables and gradients: weights and every optimizer slot after each of the recorded steps."""
    with np.load(os.path.join(GOLDEN, "reference_optimizer.npz")) as f:
        rec = _nested(f, case)
    optimizer, macro = str(rec["optimizer"]), int(rec["macro"])
    names = [str(n) for n in rec["names"]]
    assert len(names) == len(ORACLE_NAMES)
    cfg = small_config()
    cfg["optimizer"] = optimizer
    cfg.update(json.loads(str(rec["dtypes"])))
    bf16 = cfg["calculation_dtype"] == "bfloat16"
    sh = orc.Shapes(cfg)
    params = {o: rec["init"][n].astype(np.float32) for o, n in zip(ORACLE_NAMES, names)}
    state = {o: {} for o in ORACLE_NAMES}  # slots appear on first use, zero (src/optimizer/backend.py:23-25)
    opt_key = optimizer.replace(":", "_")
    for s in sorted(rec["steps"], key=int):
        step = rec["steps"][s]
        global_step = FIRST_STEP + int(s) * macro
        lr = orc.learning_rate(sh, global_step)
        assert abs(lr - float(step["lr"])) <= 1e-7 * abs(lr), (lr, float(step["lr"]))
        seen = None
        for i in range(macro):
            grads = {o: step["grads"][str(i)][str(j)].astype(np.float32) for j, o in enumerate(ORACLE_NAMES)}
            if bf16:  # the graph hands the optimizer activation-dtype gradients (src/optimizer/__init__.py:180)
                grads = {o: orc.round_bf16(g).reshape(g.shape) for o, g in grads.items()}
            last = i == macro - 1
            seen = orc.optimizer_step(sh, params, grads, state, lr, step_count=global_step + 1, seen=seen, assign=last)
        want = step["state"]
        for o, n in zip(ORACLE_NAMES, names):
            assert rel(params[o], want[n]) < 2e-6, (case, s, o)
        # every slot the reference created: "<stage><k>v/<variable>/<optimizer string>/<slot>"
        slots = {k: v for k, v in want.items() if k not in names}
        assert slots, "no optimizer slots recorded"
        seen_slots = 0
        for key, value in slots.items():
            scope, _, tail = key.partition("/")
            var_name, _, slot = tail.rpartition(f"/{opt_key}/")
            o = ORACLE_NAMES[names.index(var_name)]
            assert slot in state[o], (case, key, sorted(state[o]))
            got = np.asarray(state[o][slot], np.float32).reshape(value.shape)
            scale = max(float(np.abs(value).max()), 1e-30)
            assert float(np.abs(got - value).max()) <= 3e-6 * scale, (case, s, key)
            seen_slots += 1
        assert seen_slots == sum(len(v) for v in state.values()), "the oracle holds slots the reference does not"


def test_full_training_step_matches_the_reference_get_train_model():
    """src/run/train.py:get_train_model itself, executed (tests/golden/make_reference_vectors.py:run_training_step):
    build() -- token gather-embed, input linear, the reversible body as RevGradOps, LM head, softmax cross-entropy +
    z-loss, accuracy -- and get_optimizer -- the reverse sweep through every op's gradient(), RevGradOp.gradient's
    recompute and reconstruction, the optimizer chain -- on the oracle's seeded weights and tokens, two consecutive
    steps.  The oracle (float64 forward / backward, fp32 optimizer) must reproduce logits, loss, accuracy, every
    gradient, every updated weight and every optimizer slot; variables are matched BY CREATION ORDER, which pins that
    order too."""
    with np.load(os.path.join(GOLDEN, "reference_training_step.npz")) as f:
        rec = _nested(f, "tiny")
    cfg = small_config()
    sh = orc.Shapes(cfg)
    names = [str(n) for n in rec["oracle_names"]]
    created = [str(n) for n in rec["created"]]
    assert names == orc.param_names(sh) and len(created) == len(names)
    params = {n: rec["init"][n].astype(np.float32).copy() for n in names}
    state = orc.init_opt_state(params)
    tok_x, tok_y = rec["tok_x"], rec["tok_y"]
    opt_key = sh.optimizer.replace(":", "_")
    for s in sorted(rec["steps"], key=int):
        step = rec["steps"][s]
        gs = FIRST_STEP + int(s)
        trace = {}
        seen = {k: v.astype(np.float64) for k, v in params.items()}
        loss, acc, fwd_state = orc.forward(sh, seen, tok_x, tok_y, dtype=np.float64, trace=trace)
        assert abs(loss - float(step["loss"])) < 1e-5 * abs(loss), (loss, float(step["loss"]))
        assert abs(acc - float(step["accuracy"])) < 1e-7
        logits = trace["logits"] if "logits" in trace else fwd_state.get("logits")
        if logits is not None:
            assert rel(np.asarray(logits).reshape(step["logits"].shape), step["logits"]) < 1e-4
        grads = orc.backward(sh, fwd_state)
        for o, n in zip(names, created):
            want = step["grads"][n]
            # the recorded run computes in float32 (calculation_dtype of the config) and the reversible backward
            # differentiates at RECONSTRUCTED activations: fp32 rounding of that against the float64 oracle
            assert rel(grads[o].reshape(want.shape), want) < 2e-4, (s, o, n)
        lr = orc.learning_rate(sh, gs)
        assert abs(lr - float(step["lr"])) <= 1e-7 * lr
        orc.optimizer_step(sh, params, {k: v.astype(np.float32) for k, v in grads.items()}, state, lr, step_count=gs + 1)
        for o, n in zip(names, created):
            want = step["state"][n]
            assert rel(params[o].reshape(want.shape), want) < 1e-4, (s, o)  # fp32 run vs float64 oracle, through sm3
        n_slots = 0
        for key, value in step["state"].items():
            if key in created:
                continue
            _, _, tail = key.partition("/")
            var_name, _, slot = tail.rpartition(f"/{opt_key}/")
            o = names[created.index(var_name)]
            got = np.asarray(state[o][slot], np.float32).reshape(value.shape)
            scale = max(float(np.abs(value).max()), 1e-30)
            assert float(np.abs(got - value).max()) <= 1e-3 * scale, (s, key)
            n_slots += 1
        assert n_slots >= 2 * len(names)


def test_graph_slot_names_match_what_the_reference_creates():
    """The names of the optimizer slots in the reference's graph -- what a TF checkpoint of the reference holds --
    against obst.optimizer.backend.graph_slot_name, for every recorded optimizer string: `<stage><k>v/<variable>/
    <optimizer string>/<slot>`."""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.optimizer.backend import graph_slot_name

    class P:
        pass

    with np.load(os.path.join(GOLDEN, "reference_optimizer.npz")) as f:
        for case in OPT_CASES:
            rec = _nested(f, case)
            names = [str(n) for n in rec["names"]]
            p = P()
            p.optimizer = str(rec["optimizer"])
            recorded = sorted(k for k in rec["steps"]["0"]["state"] if k not in names)
            ranks = {n: rec["init"][n].ndim for n in names}
            opt_key = p.optimizer.replace(":", "_")
            made = []
            for key in recorded:
                _, _, tail = key.partition("/")
                var_name, _, slot = tail.rpartition(f"/{opt_key}/")
                made.append(graph_slot_name(p, var_name, slot, names.index(var_name), ranks[var_name]))
            assert sorted(made) == recorded, (case, [m for m in made if m not in recorded][:3])


@pytest.mark.parametrize("name", ["32big_mixer", "32ctx_mixer", "32mixer_group", "small"])
def test_model_parameter_derivations_match_the_reference_dataclass(name):
    """src/dataclass.py:ModelParameter run on the shipped configs (tests/golden/reference_dataclass.json) against the
    host mirror obst/dataclass.py: every derived size, dimension, dtype, embedding / optimizer string the step uses."""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.dataclass import ModelParameter, gpt_mode
    with open(os.path.join(GOLDEN, "reference_dataclass.json")) as f:
        want = json.load(f)[name]
    if name == "small":
        cfg = small_config()
    else:  # the package's own copy of the shipped config: checked here against the reference's JSON, as parsed there
        from obst.configs import mixer_config
        cfg = mixer_config(name)
    p = ModelParameter(gpt_mode(cfg))

    def dims(ds):
        return [[d.name, int(d.size)] for d in ds]

    got = {"feature_dims": dims(p.feature_dims), "intermediate": dims(p.intermediate),
           "named_dims": dims([p.head_dim, p.key_dim, p.vocab_dim, p.batch_dim, p.sequence_dim, p.token_patch_dim,
                               p.macro_batch_dim]),
           "token_dim_shape": dims(p.token_dim_shape),
           "learning_rate_config": {k: [v.start_step, v.final_step, v.factor] for k, v in p.learning_rate_config.items()},
           "block_config": [{"layer": list(b.layer), "skip": bool(b.skip)} for b in p.block_config]}
    for key, value in got.items():
        assert value == want[key], (key, value, want[key])
    dtypes = {"float32": "float32", "bfloat16": "bfloat16"}
    for key, value in want["dtypes"].items():
        mine = getattr(p, key)
        assert dtypes[value] in str(mine), (key, mine, value)
    skip = {"feature_dims", "intermediate", "named_dims", "token_dim_shape", "learning_rate_config", "block_config",
            "dtypes"}
    for key, value in want.items():
        if key in skip:
            continue
        mine = getattr(p, key)
        if isinstance(value, float):
            assert abs(mine - value) <= 1e-12 * max(1.0, abs(value)), (key, mine, value)
        else:
            assert (list(mine) if isinstance(value, list) else mine) == value, (key, mine, value)


def test_sampling_loop_matches_the_reference_autoregressive_model():
    """src/run/inference.py:autoregressive_model, executed: the forward pass per position, log(-log U) * (-temperature)
    added to the logits, argmax, shift by one, write-back at `position` -- greedy and at temperature 0.7, fed the
    recorded uniform draws."""
    with np.load(os.path.join(GOLDEN, "reference_sampling.npz")) as f:
        rec = _nested(f, "tiny")
    cfg = small_config()
    sh = orc.Shapes(cfg)
    names = [str(n) for n in rec["oracle_names"]]
    params = {n: rec["init"][n].astype(np.float64) for n in names}
    prompt = rec["prompt"]
    first = int(rec["initial_pos"])
    for r in sorted(rec["runs"], key=int):
        run = rec["runs"][r]
        uniform = run["uniform"]

        def uniform_fn(pos, uniform=uniform):
            return uniform[pos - first].reshape(prompt.shape[0], sh.sequence_length, -1)

        got = orc.sample_tokens(sh, params, prompt, first, temperature=float(run["temperature"]), uniform_fn=uniform_fn)
        want = run["tokens"].reshape(prompt.shape)
        assert np.array_equal(got[:, :first], prompt[:, :first])
        assert np.array_equal(got, want), (float(run["temperature"]), np.argwhere(got != want)[:5])


def test_learning_rate_modules_match_the_reference():
    """src/optimizer/learning_rate.py:get_learning_rate, executed, over 16 steps of three module combinations
    (linear_warmup, exponential_decay, linear_decay, lower_bound, upper_bound) against orc.learning_rate and the host
    mirror obst/optimizer/learning_rate.py."""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.optimizer.learning_rate import get_learning_rate
    with open(os.path.join(GOLDEN, "reference_learning_rate.json")) as f:
        recs = json.load(f)
    for name, rec in recs.items():
        cfg = small_config()
        cfg["learning_rate_config"] = rec["schedule"]
        sh = orc.Shapes(cfg)
        params = ModelParameter(gpt_mode(cfg))
        for step, want in zip(rec["steps"], rec["values"]):
            got = orc.learning_rate(sh, step)
            assert abs(got - want) <= 2e-7 * abs(want) + 1e-12, (name, step, got, want)
            params.global_step = step
            mine = float(get_learning_rate(params, [], []).learning_rate)
            assert abs(mine - want) <= 2e-7 * abs(want) + 1e-12, (name, step, mine, want)


def _toy_documents(kind, n_files=3, seed=9):  # the corpus of tests/golden/make_reference_vectors.py:input_documents
    rng = np.random.default_rng(seed)
    files = []
    for _ in range(n_files):
        docs = []
        for _ in range(int(rng.integers(2, 4))):
            tokens = rng.integers(32, 127, int(rng.integers(40, 100))).tolist()
            docs.append(tokens if kind == "int64" else bytes(tokens).decode("ascii"))
        files.append(docs)
    return files


@pytest.mark.parametrize("kind", ["int64", "bytes"])
def test_text_input_pipeline_matches_the_reference_gpt_neo_input(tmp_path, kind):
    """src/inputs.py:gpt_neo_input, executed on toy TFRecord files (stand-ins for tf.data.TFRecordDataset /
    interleave / window / batch / parse_single_example): the windows, their order across interleaved files, the
    file-to-slice assignment with and without filename shuffling, the x / y shift -- against obst.inputs.gpt_neo_input
    on the same files.  (The reference run ends with an error on a final partial batch; the package drops it.)"""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.inputs import gpt_neo_input, write_text_tfrecord
    for i, docs in enumerate(_toy_documents(kind)):
        count = sum(len(d) for d in docs)
        write_text_tfrecord(str(tmp_path / f"toy_{kind}_{i}_{count}.tfrecord"), docs)
    with np.load(os.path.join(GOLDEN, "reference_input_pipeline.npz")) as f:
        golden = {k: f[k] for k in f.files if k.startswith(kind + "|")}
    checked = 0
    for shuffle in (0, 1):
        for slice_index, slice_count in ((0, 1), (1, 2)):
            for interleaved in (1, 2):
                cfg = small_config()
                cfg.update(dataset_configs=[{"path": str(tmp_path / "*.tfrecord"), "type": "text", "weight": 1}],
                           interleaved_datasets=interleaved, shuffle_input_filenames=shuffle, data_seed=7,
                           use_random_dataloader=False, token_patch_size=1)
                params = ModelParameter(gpt_mode(cfg))
                batches = list(gpt_neo_input(params, 2, slice_index, slice_count))
                key = f"{kind}|shuffle{shuffle}|slice{slice_index}of{slice_count}|interleave{interleaved}"
                want_x, want_y = golden[key + "|token_x"], golden[key + "|token_y"]
                assert len(batches) == len(want_x), (key, len(batches), len(want_x))
                for b, wx, wy in zip(batches, want_x, want_y):
                    assert np.array_equal(b["token_x"].numpy(), wx), key
                    assert np.array_equal(b["token_y"].numpy(), wy), key
                checked += len(batches)
    assert checked > 40


def test_resume_simulation_matches_the_reference_simulate_data_pipeline(tmp_path):
    """src/inputs.py:simulate_data_pipeline and split_files(runs_log=...), executed on 60 random file lists / run logs,
    and gpt_neo_input resumed from a log on the toy files: which files a resumed run leaves out, how many tokens of
    each record it skips, and the windows it then reads -- against obst.inputs."""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.inputs import gpt_neo_input, simulate_data_pipeline, split_files, write_text_tfrecord
    with open(os.path.join(GOLDEN, "reference_runs_log.json")) as f:
        golden = json.load(f)
    dropped = partial = 0
    for case in golden["cases"]:
        drop, consumed = simulate_data_pipeline(case["runs_log"], sorted(case["files"]))
        assert drop == case["drop"], case
        assert consumed == case["consumed"], case
        for i, (want_files, want_skips) in enumerate(case["split"]):
            files, skips = split_files(list(case["files"]), i, case["slice_count"], case["seed"], case["runs_log"])
            assert list(files) == want_files and list(skips) == want_skips, case
        dropped += sum(drop)
        partial += sum(1 for d, c in zip(drop, consumed) if c and not d)
    assert dropped > 50 and partial > 20  # the vectors exercise both outcomes

    for i, docs in enumerate(_toy_documents("int64")):
        write_text_tfrecord(str(tmp_path / f"toy_int64_{i}_{sum(len(d) for d in docs)}.tfrecord"), docs)
    for key, rec in golden["resumed"].items():
        interleaved = int(key.split("|")[0][len("interleave"):])
        cfg = small_config()
        cfg.update(dataset_configs=[{"path": str(tmp_path / "*.tfrecord"), "type": "text", "weight": 1}],
                   interleaved_datasets=interleaved, shuffle_input_filenames=0, data_seed=7,
                   use_random_dataloader=False, token_patch_size=1)
        batches = list(gpt_neo_input(ModelParameter(gpt_mode(cfg)), 2, 0, 1, rec["runs_log"]))
        assert len(batches) == len(rec["token_x"]) > 0, key
        for b, want in zip(batches, rec["token_x"]):
            assert np.array_equal(b["token_x"].numpy(), np.asarray(want)), key
