"""Golden vectors from the REFERENCE'S OWN PYTHON, run in the build container (where /root/reference exists).

    python tests/golden/make_reference_vectors.py        ->  tests/golden/reference_*.npz

`src.*` of /root/reference is imported unmodified; its two third-party tensor libraries (tensorflow 2.7,
mesh-tensorflow: not installable here) are replaced by the eager NumPy stand-ins of oracle/ref_shim (see its
README).  Recorded:

  reference_optimizer.npz      src/optimizer/__init__.py:get_optimizer itself -- learning-rate context, step / beta /
                               step_count arithmetic, every stage of the optimizer string, the weight-decay rule and
                               the assignment -- driven by a loss whose gradient is a given array per variable (a
                               test-side mtf.Operation: <var, G>; the reference's reverse sweep calls its gradient()),
                               for several optimizer strings, three consecutive steps each, and the macro-batching
                               ("add" then "update") call pair of src/run/train.py:35-74
  reference_learning_rate.npz  src/optimizer/learning_rate.py:get_learning_rate over steps and module combinations
  reference_layers.npz         forward values of src/model/normalization.py:norm, the activations of
                               src/model/activation.py, src/mtf_wrapper.py:softmax_cross_entropy_with_logits
  reference_dataclass.json     derived sizes of src/dataclass.py:ModelParameter for the shipped configs

tests/test_oracle_pins.py compares oracle/obst_oracle.py (and the host mirror's ModelParameter) with these files.
"""
import copy
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.environ.get("OBST_GOLDEN_OUT", HERE)  # where the vectors go (tests/test_golden_regeneration.py: a scratch directory)
REFERENCE = "/root/reference"
sys.path.insert(0, os.path.join(ROOT, "oracle", "ref_shim"))
sys.path.insert(0, REFERENCE)
sys.path.insert(1, ROOT)

import mesh_tensorflow as mtf  # noqa: E402  (the stand-in)
import tensorflow as tf  # noqa: E402  (the stand-in)

from src import utils_core  # noqa: E402
from src.dataclass import BlockArgs, ModelParameter  # noqa: E402
from src.optimizer import get_optimizer  # noqa: E402
from src.optimizer.learning_rate import get_learning_rate  # noqa: E402
from src.utils_mtf import get_variable  # noqa: E402

from oracle import obst_oracle as orc  # noqa: E402  (only for the shared tiny config and seeded arrays)


class LinearLoss(mtf.Operation):
    """loss = sum_v <v, G_v>: its gradient hands every variable the array G_v (times the incoming gradient)."""

    def __init__(self, values, grads):
        super().__init__(values, name="linear_loss")
        self.grads = grads
        total = sum(float((v.np.astype(np.float64) * g).sum()) for v, g in zip(values, grads))
        self._outputs = [mtf.Tensor(self, [], tf.float32, value=np.float32(total))]

    def gradient(self, grad_ys):
        dy = grad_ys[0]
        # a gradient reaches get_optimizer in the dtype of the tensor it belongs to (the activation dtype)
        return [mtf.cast(mtf.import_fully_replicated(v.mesh, g.astype(np.float32), v.shape, "given_gradient"), v.dtype)
                * mtf.cast(dy, v.dtype) for v, g in zip(self.inputs, self.grads)]


def fresh_params(config):
    utils_core.NAME_INDICES.clear()
    tf.compat.v1.get_variable_scope().stack.clear()
    params = ModelParameter(copy.deepcopy(config))
    params.mesh = mtf.Mesh(mtf.Graph(), "mesh")
    return params


class Given:
    """Initializer returning a fixed array."""

    def __init__(self, array):
        self.array = array

    def __call__(self, shape, dtype=None):
        return self.array.reshape(shape)


def make_variables(params, specs):
    """specs: [(scope, dims, array)]; created through the reference's get_variable under that variable scope."""
    out = []
    for scope, dims, array in specs:
        with tf.compat.v1.variable_scope(scope):
            t = get_variable(params, "get_variable0", dims, Given(array), True, params.variable_dtype)
        out.append(t)
    return out


def variable_specs(params, rng):
    """A cross-section of the mixer's variables: scope names decide the weight-decay rule
    (src/optimizer/__init__.py:49-57), ranks decide sm3 / adam."""
    H, K = params.head_dim, params.key_dim
    inter = params.intermediate[0]
    seq = params.sequence_dim
    aseq = mtf.Dimension("_" + seq.name, seq.size)

    def arr(dims, scale=1.0):
        return (rng.standard_normal([d.size for d in dims]) * scale).astype(np.float32)

    specs = [
        ("gpt0v/body0v/0_0v/bottleneck_group_linear_0v/activated_linear_in_0v/linear_0v/orthogonal_var_0v", [H, K, inter], 0.05),
        ("gpt0v/body0v/0_0v/bottleneck_group_linear_0v/activated_linear_out_0v/linear_0v/orthogonal_var_0v", [H, inter, K], 0.05),
        ("gpt0v/body0v/0_0v/norm_0v/normal_var_0v", [H, K], 1.0),
        ("gpt0v/body0v/0_1v/attention_0v/embed_0v/normal_var_0v", [H, seq, aseq], 0.02),
        ("gpt0v/input0v/linear_to_features_0v/orthogonal_var_0v", [inter, H, K], 0.05),
        ("gpt0v/body0v/0_0v/rezero_0v/get_var_0v", [], 0.3),
    ]
    return [(scope, dims, arr(dims, s)) for scope, dims, s in specs]


def run_optimizer_case(config, optimizer, steps, macro=1, seed=0):
    cfg = copy.deepcopy(config)
    cfg["optimizer"] = optimizer
    cfg["macro_batching"] = macro
    params = fresh_params(cfg)
    rng = np.random.default_rng(seed)
    specs = variable_specs(params, rng)
    values = make_variables(params, specs)
    names = list(params.variable_cache.keys())
    record = {"names": names, "steps": []}
    record["init"] = {n: params.variable_cache[n].master.copy() for n in names}
    for step in range(steps):
        tf.GLOBAL_STEP[0] = np.asarray(cfg.get("_first_step", 1500) + step * macro, np.int64)
        grads_per_slice = []
        idx = mtf.constant(params.mesh, 0, [], params.optimizer_calculation_dtype)
        update_ops = []
        for i in range(1, macro + 1):
            utils_core.NAME_INDICES.clear()
            grads = [(rng.standard_normal(v.shape.to_integer_list) * 10 ** rng.uniform(-3, -1)).astype(np.float32)
                     for v in values]
            grads_per_slice.append(grads)
            # what the graph of this slice reads: the cache entry (a Variable, or the chained tensor's op)
            seen = [params.variable_cache[n].outputs[0] for n in names]
            loss = LinearLoss(seen, grads).outputs[0]
            last = i == macro
            update_ops, lr = get_optimizer([loss], params, idx, "update" if last else "add")
            if not last:  # src/run/train.py:63-69
                idx += 1
                for tensor in update_ops:
                    from src.mtf_wrapper import get_variable_for_tensor
                    op = get_variable_for_tensor(tensor)
                    chained = mtf.stop_gradient(mtf.cast(tensor, op.activation_dtype))
                    params.variable_cache[op.full_name] = chained.operation
        for op in update_ops:  # every new value from the OLD variable values, then the assignments
            op.stage()
        for op in update_ops:
            op.apply()
        # Within one get_train_model graph the cache entries of variables AND slots are replaced by the chained
        # tensors' ops (src/run/train.py:63-69).  The reference builds that graph once and re-runs it every step, so
        # the next step's first slice reads the variables again: put them back.
        for var in [v for v in params.mesh.graph.operations if isinstance(v, mtf.Variable)]:
            if hasattr(var, "full_name"):
                params.variable_cache[var.full_name] = var
        state = {v.full_name: v.master.copy() for v in params.mesh.graph.operations
                 if isinstance(v, mtf.Variable) and hasattr(v, "full_name")}
        record["steps"].append({"grads": grads_per_slice, "lr": float(lr.np), "state": state})
        params.mesh.graph.operations[:] = [op for op in params.mesh.graph.operations if isinstance(op, mtf.Variable)]
    return record


def run_training_step(config, steps=2, seed=3, first_step=1500):
    """src/run/train.py:get_train_model itself -- build() (input embedding, the reversible body of RevGradOps, LM
    head, softmax cross-entropy + z-loss, accuracy) and get_optimizer (the reverse sweep through every op's
    gradient(), RevGradOp.gradient's recompute / reconstruction, the optimizer chain) -- on the oracle's seeded
    weights and tokens.  Weights are imposed in creation order (that order is part of what is pinned)."""
    import src.optimizer as ref_optimizer
    from src.run.train import get_train_model
    params = fresh_params(config)
    params.train = True
    sh = orc.Shapes(config)
    weights = orc.init_params(sh, seed=seed)
    names = orc.param_names(sh)
    tok_x, tok_y = orc.synthetic_tokens(sh)
    record = {"oracle_names": names, "tok_x": tok_x, "tok_y": tok_y, "steps": []}
    record["init"] = {n: weights[n] for n in names}
    grads = {}
    original_update = ref_optimizer.update

    def recording_update(ctx):
        grads[ctx.var.full_name] = ctx.grad.np.copy()
        return original_update(ctx)

    ref_optimizer.update = recording_update
    try:
        for step in range(steps):
            utils_core.NAME_INDICES.clear()
            tf.GLOBAL_STEP[0] = np.asarray(first_step + step, np.int64)
            if step == 0:
                mtf.VARIABLE_VALUES[:] = [weights[n] for n in names]
            shape = params.token_dim_shape
            tx = mtf.import_fully_replicated(params.mesh, tok_x.reshape(shape.to_integer_list).astype(np.int32), shape, "x")
            ty = mtf.import_fully_replicated(params.mesh, tok_y.reshape(shape.to_integer_list).astype(np.int32), shape, "y")
            grads.clear()
            out = get_train_model(params, None, None, None, tx, ty, None, None, None, None)
            _, token_out, lr, total_loss, _, _, _, token_loss, accuracy, update_ops, _ = out
            assert not mtf.VARIABLE_VALUES, "fewer trainable variables were created than the oracle names"
            created = [n for n, v in params.variable_cache.items() if isinstance(v, mtf.Variable) and v.trainable]
            rec = {"logits": token_out.np.copy(), "loss": float(total_loss.np), "token_loss": float(token_loss.np),
                   "accuracy": float(accuracy.np), "lr": float(lr.np),
                   "grads": {n: grads[n] for n in created}}
            for op in update_ops:
                op.stage()
            for op in update_ops:
                op.apply()
            rec["state"] = {v.full_name: v.master.copy() for v in params.mesh.graph.operations
                            if isinstance(v, mtf.Variable) and hasattr(v, "full_name")}
            record["created"] = created
            record["steps"].append(rec)
            # the reference builds the graph once and re-runs it; here the step is rebuilt on the same variables
            params.mesh.graph.operations[:] = [op for op in params.mesh.graph.operations if isinstance(op, mtf.Variable)]
            for var in params.mesh.graph.operations:
                if hasattr(var, "full_name"):
                    params.variable_cache[var.full_name] = var
            params.cached_parameters.clear()
    finally:
        ref_optimizer.update = original_update
    return record


def run_model_case(config, steps=1, first_step=1500, seed=11):
    """get_train_model on an arbitrary config with the reference's OWN initialisers (orthogonal / normal / zeros --
    drawn from NumPy's generator, see oracle/ref_shim).  Records the created trainable variables in creation order,
    the tokens, and per step: logits, loss, accuracy, lr, every gradient, every updated variable.  Consumed by the
    product's GPU tests, which import the weights in creation order."""
    import random

    import src.optimizer as ref_optimizer
    from src.run.train import get_train_model
    random.seed(seed)  # OrthogonalInit draws its seed from `random` (src/model/backend.py:23)
    tf._SEEDS[0] = 1000 * seed
    params = fresh_params(config)
    params.train = True
    rng = np.random.default_rng(seed)
    macro = params.macro_batching
    shape = mtf.Shape(params.token_dim_shape.dims)
    ints = shape.to_integer_list
    ints[0] *= macro
    full = mtf.Shape([mtf.Dimension(shape.dims[0].name, ints[0])] + shape.dims[1:])
    window = rng.integers(0, params.vocab_size, [ints[0], params.sequence_length + 1])
    tok_x, tok_y = window[:, :-1].astype(np.int32), window[:, 1:].astype(np.int32)
    record = {"config": json.dumps(config), "tok_x": tok_x, "tok_y": tok_y, "steps": []}
    grads = {}
    original_update = ref_optimizer.update

    def recording_update(ctx):
        grads.setdefault(ctx.var.full_name, []).append(ctx.grad.np.copy())
        return original_update(ctx)

    ref_optimizer.update = recording_update
    try:
        for step in range(steps):
            utils_core.NAME_INDICES.clear()
            tf.GLOBAL_STEP[0] = np.asarray(first_step + step * macro, np.int64)
            tx = mtf.import_fully_replicated(params.mesh, tok_x.reshape(ints), full, "x")
            ty = mtf.import_fully_replicated(params.mesh, tok_y.reshape(ints), full, "y")
            grads.clear()
            out = get_train_model(params, None, None, None, tx, ty, None, None, None, None)
            _, token_out, lr, total_loss, first_loss, last_loss, _, token_loss, accuracy, update_ops, _ = out
            # (with macro-batching src/run/train.py:57-60 lists the variables twice; it deduplicates a private copy)
            variables = list(dict.fromkeys(v for v in params.mesh.graph.operations
                                           if isinstance(v, mtf.Variable) and hasattr(v, "full_name")))
            created = [v.full_name for v in variables if v.trainable]
            if step == 0:
                record["created"] = created
                record["shapes"] = {v.full_name: np.asarray(v.shape.to_integer_list, np.int64) for v in variables if v.trainable}
                record["init"] = {v.full_name: v.master.copy() for v in variables if v.trainable}
            rec = {"logits": token_out.np.copy(), "loss": float(total_loss.np), "first_loss": float(first_loss.np),
                   "last_loss": float(last_loss.np), "accuracy": float(accuracy.np), "lr": float(lr.np),
                   "grads": {n: grads[n][-1] for n in created if n in grads}}
            for op in update_ops:
                op.stage()
            for op in update_ops:
                op.apply()
            rec["state"] = {v.full_name: v.master.copy() for v in variables if v.trainable}
            record["steps"].append(rec)
            params.mesh.graph.operations[:] = variables
            for var in variables:
                params.variable_cache[var.full_name] = var
            params.cached_parameters.clear()
    finally:
        ref_optimizer.update = original_update
    return record


def model_cases():
    """Configs of the layer zoo: every registered layer / strategy the product implements, at toy sizes."""
    base = small_config()
    base.update(depth=1)

    def cfg(**kw):
        c = copy.deepcopy(base)
        c.update(kw)
        return c

    norm = "norm-shift-scale-features-group"
    mix = "attention-biased_attention_map-absolute-input_as_value-shared"

    def blocks(*layer_lists, **extra):
        return [dict({"layer": list(ls)}, **extra) for ls in layer_lists]

    def one(layers, **kw):  # a single non-reversible block: every gradient is first-hand, no reconstruction noise
        return cfg(memory_reduction_strategy="none", block_config=blocks([norm] + list(layers)), **kw)

    cases = {
        "mixer_depth2": cfg(depth=2),
        "strategy_none": cfg(memory_reduction_strategy="none"),
        # (`shared` variables cannot be combined with 'checkpoint' in the reference: recompute_grad's scope shifts the
        # block name src/model/backend.py:59-62 parses, and the shared-variable counter divides by zero)
        "strategy_checkpoint_skip": cfg(memory_reduction_strategy="checkpoint",
                                        block_config=blocks([norm, "feed_forward-in:relu-in:norm-in:shift-in:scale-in:features"],
                                                            [norm, mix.replace("-shared", "")], skip=True)),
        "strategy_none_skip": cfg(memory_reduction_strategy="none",
                                  block_config=blocks([norm, "feed_forward-in:relu"], [norm, mix], skip=True)),
        "feed_forward_glu": cfg(block_config=blocks([norm, "feed_forward-in:mish-in:glu-in:norm-in:shift-in:scale-in:features"],
                                                    [norm, "feed_forward-in:silu-in:glu_add"])),
        "activations": cfg(memory_reduction_strategy="none",
                           block_config=blocks([norm, "feed_forward-in:lecun_tanh"], [norm, "feed_forward-in:softsign"],
                                               [norm, "feed_forward-in:mtf_mish"], [norm, "feed_forward-in:sigmoid-in:tanh"],
                                               [norm, "feed_forward-in:gelu"])),
        # ('exp' is registered as ACTIVATIONS['exp'] = mtf.exp and called with the BlockArgs, not its tensor,
        # src/model/activation.py:201,211: it cannot run in the reference)
        "norm_variants": cfg(memory_reduction_strategy="none", block_config=blocks(
            ["norm", "feed_forward-in:relu"], ["norm-scale", mix], ["norm-shift-group", "feed_forward-in:relu"],
            ["norm-scale-shift-features", mix])),
        "bottleneck_variants": cfg(memory_reduction_strategy="none", block_config=blocks(
            [norm, "bottleneck_group_linear-in:gelu-in:norm-in:scale-mid:silu-mid:glu"],
            [norm, "bottleneck_group_linear-in:relu-mid:relu-mid:norm-mid:shift-mid:scale-mid:features-out:tanh"])),
        "standalone_activations": cfg(memory_reduction_strategy="none", block_config=blocks(
            [norm, "activation-silu", "feed_forward-in:relu"], [norm, "activation-mish", mix],
            [norm, "activation-lecun_tanh", "activation-softsign", mix])),
        "cumsum_cummean": cfg(block_config=blocks([norm, "cumsum"], [norm, "cummean", "rezero"])),
        # (reduced_half_linear reshapes a [.., heads, _features] product back to the head-reduced input's shape,
        # src/model/basic.py:72-74,89-90: with more than one head mtf.reshape refuses the size change, so the layer has
        # no reference behaviour here; sum_heads alone leaves a head-less tensor for the residual add)
        "group_linear": cfg(block_config=blocks([norm, "group_linear"], [norm, mix])),
        "transpose_sequence_features": cfg(sequence_length=8, block_config=blocks(
            [norm, "transpose_sequence_features", "group_linear", "transpose_sequence_features"], [norm, mix])),
        "attention_embedded": one(["attention-dot_product-embedded-absolute-in:relu"]),
        "attention_context_biased_softmax": one(["attention-dot_product-context-input_as_value-biased_softmax-absolute"]),
        "attention_maps": one(["attention-biased_softmax-scale_attention_map-biased_attention_map-input_as_value-absolute"]),
        "attention_shared_key_value": one(["attention-dot_product-embedded-shared_key_value-absolute"]),
        "attention_shared_key_value_maps": one(
            ["attention-dot_product-context-shared_key_value-biased_attention_map-scale_attention_map-absolute"]),
        # the sinusoid's schedule is exp(feature index): beyond a few features its sine is rounding noise in fp32
        # (src/model/embedding.py:167-169), so these run with 2 heads x 4 features
        "attention_relative": one(["attention-dot_product-positional-relative-input_as_value"],
                                  position_embedding="relative", heads=2, features_per_head=4, vocab_weight_factorization=0.5),
        "attention_relative_learned": one(["attention-dot_product-positional-relative-learned-input_as_value"],
                                          position_embedding="relative", heads=2, features_per_head=4,
                                          vocab_weight_factorization=0.5),
        "attention_axial": one(["attention-dot_product-positional-axial-input_as_value"]),
        "attention_stack_revnet": cfg(block_config=blocks(
            [norm, "attention-dot_product-embedded-absolute-in:relu"],
            [norm, "attention-dot_product-context-input_as_value-biased_softmax-absolute"],
            [norm, "attention-dot_product-embedded-shared_key_value-absolute"])),
        "split_path": cfg(block_config=blocks([norm, f"split_path-add;feed_forward-in:relu;{mix}"],
                                              [norm, f"split_path-multiply;feed_forward-in:tanh;{mix}"])),
        "initial_position_embedding": cfg(use_initial_position_embedding=True, position_embedding="axial"),
        "macro_batching": cfg(macro_batching=2, depth=2),
        # bf16 activations over fp32 slices, as the 32*_mixer configs run: the stand-ins round every op's output to
        # bfloat16 -- a model of WHERE the graph rounds (each MTF op), not of XLA's internal accumulation
        "mixer_bf16_none": cfg(calculation_dtype="bfloat16", memory_reduction_strategy="none", embedding_stddev=0.04, vocab_weight_factorization=1.0,
                               train_batch_size=4),
        "mixer_bf16_revnet": cfg(calculation_dtype="bfloat16", embedding_stddev=0.04, vocab_weight_factorization=1.0,
                               train_batch_size=4),
        # the same two models (same seeds, hence the same weights and tokens) in float32: the yardstick for how far
        # bf16 takes ANY implementation from the truth
        "mixer_f32_none_twin": cfg(memory_reduction_strategy="none", embedding_stddev=0.04, vocab_weight_factorization=1.0,
                               train_batch_size=4),
        "mixer_f32_revnet_twin": cfg(embedding_stddev=0.04, vocab_weight_factorization=1.0,
                               train_batch_size=4),
        "adam_chain": cfg(optimizer="adam-learning_rate", depth=2),
    }
    return cases


def run_sampling(config, seed=5, initial_pos=4, temperatures=(0.0, 0.7)):
    """src/run/inference.py:autoregressive_model (language branch) on the oracle's seeded weights: the whole sampling
    loop -- forward pass per position, Gumbel-temperature argmax, shift, write-back.  The uniform draws are recorded so
    that a checker can feed the same ones."""
    from src.run.inference import autoregressive_model
    sh = orc.Shapes(config)
    weights = orc.init_params(sh, seed=seed)
    names = orc.param_names(sh)
    rng = np.random.default_rng(seed)
    prompt = rng.integers(0, sh.vocab_size, [sh.train_batch_size, sh.sequence_length]).astype(np.int32)
    record = {"oracle_names": names, "init": {n: weights[n] for n in names}, "prompt": prompt,
              "initial_pos": initial_pos, "runs": []}
    for temperature in temperatures:
        cfg = dict(config, sampling_temperature=temperature, initial_autoregressive_position=initial_pos,
                   use_autoregressive_sampling=True)
        params = fresh_params(cfg)
        params.train = False
        mtf.VARIABLE_VALUES[:] = [weights[n] for n in names]
        draws = []

        def uniform(shape, lo, hi, draws=draws, rng=np.random.default_rng(100 + int(temperature * 10))):
            u = rng.uniform(lo, hi, shape).astype(np.float32)
            draws.append(u)
            return u

        mtf.UNIFORM_SOURCE[0] = uniform
        names_at_loop = {}

        def rewind(params=params):  # the loop body is traced once in the library: same names on every re-run
            if not names_at_loop:
                names_at_loop.update(utils_core.NAME_INDICES)
            utils_core.NAME_INDICES.clear()
            utils_core.NAME_INDICES.update(names_at_loop)
            params.cached_parameters.clear()
            params.attention_idx = 0

        mtf.BEFORE_LOOP_BODY[0] = rewind
        shape = params.token_dim_shape
        tx = mtf.import_fully_replicated(params.mesh, prompt.reshape(shape.to_integer_list), shape, "x")
        ty = mtf.import_fully_replicated(params.mesh, prompt.reshape(shape.to_integer_list), shape, "y")
        try:
            token_out, _ = autoregressive_model(params, token_x_input=tx, token_y_input=ty)
            tokens = token_out.np.copy()
        finally:
            mtf.UNIFORM_SOURCE[0] = None
            mtf.BEFORE_LOOP_BODY[0] = None
            # only the first build creates variables; later iterations must not consume imposed values
            mtf.VARIABLE_VALUES[:] = []
        record["runs"].append({"temperature": temperature, "tokens": tokens, "uniform": np.stack(draws)})
    return record


def input_documents(kind, n_files=3, seed=9):
    """Deterministic toy corpora: per file a few documents of 40-100 tokens (ints < 200, or ASCII text)."""
    rng = np.random.default_rng(seed)
    files = []
    for _ in range(n_files):
        docs = []
        for _ in range(int(rng.integers(2, 4))):
            tokens = rng.integers(32, 127, int(rng.integers(40, 100))).tolist()
            docs.append(tokens if kind == "int64" else bytes(tokens).decode("ascii"))
        files.append(docs)
    return files


def write_input_files(directory, kind):
    """TFRecord files named as the reference expects (`int64` in the name selects decode_intstring; the trailing
    number is the file's element count, src/inputs.py:34), written with the package's writer."""
    sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
    from obst.inputs import write_text_tfrecord
    paths = []
    for i, docs in enumerate(input_documents(kind)):
        count = sum(len(d) for d in docs)
        path = os.path.join(directory, f"toy_{kind}_{i}_{count}.tfrecord")
        write_text_tfrecord(path, docs)
        paths.append(path)
    return paths


def run_input_pipeline():
    """src/inputs.py:gpt_neo_input on toy TFRecord files, through stand-ins for tf.data / TFRecordDataset /
    parse_single_example: which windows come out, in which order, for which slice."""
    import tempfile

    from src.inputs import gpt_neo_input
    out = {}
    for kind in ("int64", "bytes"):
        with tempfile.TemporaryDirectory() as directory:
            write_input_files(directory, kind)
            for shuffle in (0, 1):
                for slice_index, slice_count in ((0, 1), (1, 2)):
                    for interleaved in (1, 2):
                        cfg = small_config()
                        cfg.update(dataset_configs=[{"path": os.path.join(directory, "*.tfrecord"), "type": "text",
                                                     "weight": 1}],
                                   interleaved_datasets=interleaved, shuffle_input_filenames=shuffle, data_seed=7,
                                   use_random_dataloader=False, token_patch_size=1)
                        params = fresh_params(cfg)
                        batches, partial = [], 0
                        it = iter(gpt_neo_input(params, 2, slice_index, slice_count))
                        while True:
                            try:
                                batches.append(next(it))
                            except StopIteration:
                                break
                            except ValueError:
                                # dset.batch(sub_batch_size) keeps a final partial batch, which _memory_func's
                                # fixed-shape reshape rejects (src/inputs.py:543-545): the reference run ends there
                                partial = 1
                                break
                        key = f"{kind}|shuffle{shuffle}|slice{slice_index}of{slice_count}|interleave{interleaved}"
                        out[key + "|ended_on_partial_batch"] = np.asarray(partial)
                        out[key + "|token_x"] = np.stack([np.asarray(b[0]) for b in batches])
                        out[key + "|token_y"] = np.stack([np.asarray(b[1]) for b in batches])
    return out


def run_resume_simulation(n_cases=60, seed=21):
    """src/inputs.py:simulate_data_pipeline / split_files on random file lists and run logs (pure integer host logic;
    the file names carry the token counts), plus gpt_neo_input resumed from a log on the toy files."""
    import tempfile

    from src.inputs import gpt_neo_input, simulate_data_pipeline, split_files
    rng = np.random.default_rng(seed)
    cases = []
    for _ in range(n_cases):
        n_files = int(rng.integers(1, 13))
        ctx = int(rng.choice([8, 16, 32]))
        files = [f"data/set_{i:03d}_{int(rng.integers(0, 12 * ctx))}.tfrecord" for i in range(n_files)]
        runs = []
        for _ in range(int(rng.integers(1, 4))):
            slice_count = int(rng.choice([1, 2, 4]))
            runs.append({"steps": int(rng.integers(0, 9)), "grad_accumulation": int(rng.integers(1, 3)),
                         "batch_size": slice_count * int(rng.integers(1, 4)), "slice_count": slice_count, "ctx": ctx,
                         "interleave_size": int(rng.integers(1, 4)), "token_patch_size": int(rng.choice([1, 1, 2]))})
        drop, consumed = simulate_data_pipeline(runs, sorted(files))
        shuffle_seed = int(rng.choice([0, 5]))
        slice_count = int(rng.choice([1, 2]))
        split = [split_files(list(files), i, slice_count, shuffle_seed, runs) for i in range(slice_count)]
        cases.append({"files": files, "runs_log": runs, "drop": [bool(d) for d in drop],
                      "consumed": [int(c) for c in consumed], "seed": shuffle_seed, "slice_count": slice_count,
                      "split": [[list(f), [int(k) for k in sk]] for f, sk in split]})
    resumed = {}
    with tempfile.TemporaryDirectory() as directory:
        write_input_files(directory, "int64")
        for interleaved, steps in ((1, 1), (2, 2), (2, 3)):
            cfg = small_config()
            cfg.update(dataset_configs=[{"path": os.path.join(directory, "*.tfrecord"), "type": "text", "weight": 1}],
                       interleaved_datasets=interleaved, shuffle_input_filenames=0, data_seed=7,
                       use_random_dataloader=False, token_patch_size=1)
            params = fresh_params(cfg)
            log = [{"steps": steps, "grad_accumulation": 1, "batch_size": 2, "slice_count": 1,
                    "ctx": params.sequence_length, "interleave_size": interleaved, "token_patch_size": 1}]
            batches = []
            it = iter(gpt_neo_input(params, 2, 0, 1, log))
            while True:
                try:
                    batches.append(next(it))
                except (StopIteration, ValueError):  # ValueError: the final partial batch, as in run_input_pipeline
                    break
            resumed[f"interleave{interleaved}|steps{steps}"] = {
                "runs_log": log, "token_x": [np.asarray(b[0]).tolist() for b in batches]}
    return {"cases": cases, "resumed": resumed}


def write_resume_simulation():
    rec = run_resume_simulation()
    with open(os.path.join(OUT, "reference_runs_log.json"), "w") as f:
        json.dump(rec, f)
    print("runs_log ok:", len(rec["cases"]), "cases;", {k: len(v["token_x"]) for k, v in rec["resumed"].items()})


def describe_parameters(params):
    """The derived attributes of a ModelParameter as plain JSON."""
    def dims(ds):
        return [[d.name, int(d.size)] for d in ds]

    out = {k: getattr(params, k) for k in (
        "features", "features_per_head", "heads", "depth", "sequence_length", "train_batch_size", "vocab_size",
        "group_linear_factor", "intermediate_feed_forward_multiplier", "mesh_shape", "layout", "macro_batching",
        "language_token_per_frame", "time_patch_size", "token_patch_size", "language_token_patch", "weight_decay",
        "learning_rate", "optimizer", "memory_reduction_strategy", "use_video", "use_language", "z_loss",
        "embedding_stddev", "scale_by_depth", "weight_centralisation", "weight_standardisation",
        "position_embedding", "token_embedding", "output_embedding", "masked_attention_dimensions", "rezero_lr_multiplier",
        "vocab_weight_factorization", "opt_beta1", "opt_beta2", "grad_accumulation", "input_dropout")}
    out["feature_dims"] = dims(params.feature_dims)
    out["intermediate"] = dims(params.intermediate)
    out["named_dims"] = dims([params.head_dim, params.key_dim, params.vocab_dim, params.batch_dim, params.sequence_dim,
                              params.token_patch_dim, params.macro_batch_dim])
    out["token_dim_shape"] = dims(params.token_dim_shape.dims)
    out["dtypes"] = {k: getattr(params, k).name for k in ("slice_dtype", "storage_dtype", "calculation_dtype",
                                                          "optimizer_slice_dtype", "optimizer_calculation_dtype")}
    out["learning_rate_config"] = {k: [v.start_step, v.final_step, v.factor] for k, v in params.learning_rate_config.items()}
    out["block_config"] = [{"layer": list(b.layer), "skip": bool(b.skip)} for b in params.block_config]
    return out


def flatten(prefix, obj, out):
    if isinstance(obj, dict):
        for k, v in obj.items():
            flatten(f"{prefix}|{k}", v, out)
    elif isinstance(obj, (list, tuple)) and obj and not isinstance(obj[0], str):
        for i, v in enumerate(obj):
            flatten(f"{prefix}|{i}", v, out)
    elif isinstance(obj, (list, tuple)):
        out[prefix] = np.asarray(obj)
    else:
        out[prefix] = np.asarray(obj)


def small_config():
    """The oracle's tiny mixer config, shrunk so that the committed vectors stay small."""
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(sequence_length=16, features_per_head=8, heads=4, train_batch_size=2, vocab_size=32)
    cfg.update(use_video=False, use_language=True)  # the model_mode == 'gpt' override of src/main.py:87-93
    return cfg


def main():
    if sys.argv[1:] == ["runs_log"]:  # that file alone
        write_resume_simulation()
        return
    write_resume_simulation()
    base = small_config()
    out = {}
    cases = {
        "hot_path": ("adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", 3, 1),
        "no_clip_heavyball": ("sm3-momentum:0.9:1:0-learning_rate", 2, 1),
        "adam": ("adam-learning_rate", 3, 1),
        "novograd": ("novograd-momentum:0.9:1:0-learning_rate", 2, 1),
        "clips": ("l2norm_clip:0.5-value_clip:0.01-gradient_centralisation-weight_centralisation-sm3-learning_rate", 2, 1),
        "global_clip": ("global_l2norm_clip:0.2-momentum:0.95:0.5:1-learning_rate", 2, 1),
        "graft": ("graft:adam-learning_rate", 2, 1),
        "macro2": ("adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", 2, 2),
        # the dtype set of the 32*_mixer configs: bf16 activations and storage, bf16 optimizer slots, fp32 slices
        "hot_path_bf16": ("adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", 3, 1),
        "adam_bf16": ("adam-learning_rate", 2, 1),
    }
    for name, (optimizer, steps, macro) in cases.items():
        cfg_case = dict(base)
        if name.endswith("_bf16"):
            cfg_case.update(calculation_dtype="bfloat16", storage_dtype="bfloat16", optimizer_slice_dtype="bfloat16")
        rec = run_optimizer_case(cfg_case, optimizer, steps, macro, seed=100 + list(cases).index(name))
        rec["dtypes"] = json.dumps({k: cfg_case[k] for k in ("calculation_dtype", "storage_dtype",
                                                              "optimizer_slice_dtype", "slice_dtype")})
        rec["optimizer"] = optimizer
        rec["macro"] = macro
        flatten(name, rec, out)
        print("optimizer case", name, "ok:", len(rec["steps"]), "steps")
    np.savez_compressed(os.path.join(OUT, "reference_optimizer.npz"), **out)

    out = {}
    rec = run_training_step(small_config())
    flatten("tiny", rec, out)
    print("training step ok: loss", [s["loss"] for s in rec["steps"]])
    np.savez_compressed(os.path.join(OUT, "reference_training_step.npz"), **out)

    out = {}
    failed = {}
    for name, config in model_cases().items():
        try:
            rec = run_model_case(config, steps=2 if name in ("mixer_depth2", "macro_batching", "adam_chain") else 1)
        except Exception as exc:  # noqa: BLE001 -- reported, the other cases are still recorded
            failed[name] = f"{type(exc).__name__}: {exc}"
            print("model case", name, "FAILED:", failed[name])
            continue
        flatten(name, rec, out)
        print("model case", name, "ok: loss", [round(s["loss"], 5) for s in rec["steps"]], len(rec["created"]), "variables")
    np.savez_compressed(os.path.join(OUT, "reference_models.npz"), **out)
    if failed:
        print("FAILED CASES:", json.dumps(failed, indent=1))

    out = {}
    rec = run_sampling(small_config())
    flatten("tiny", rec, out)
    np.savez_compressed(os.path.join(OUT, "reference_sampling.npz"), **out)
    print("sampling ok:", [r["tokens"].reshape(rec["prompt"].shape)[0, :10].tolist() for r in rec["runs"]])

    pipeline = run_input_pipeline()
    np.savez_compressed(os.path.join(OUT, "reference_input_pipeline.npz"), **pipeline)
    print("input pipeline ok:", {k: v.shape for k, v in list(pipeline.items())[:4]})

    # ---- src/optimizer/learning_rate.py:get_learning_rate over steps and module combinations ------------------------
    schedules = {
        "warmup": {"linear_warmup": {"final_step": 4096}},
        "warmup_exp": {"linear_warmup": {"final_step": 100}, "exponential_decay": {"start_step": 50, "factor": 0.999}},
        "warmup_linear_bounds": {"linear_warmup": {"final_step": 10}, "linear_decay": {"start_step": 20, "final_step": 400},
                                 "lower_bound": {"factor": 1e-4}, "upper_bound": {"factor": 5e-3}},
    }
    steps = [0, 1, 5, 10, 49, 50, 51, 99, 100, 101, 399, 400, 1000, 4095, 4096, 100000]
    lr_out = {}
    for name, schedule in schedules.items():
        cfg = small_config()
        cfg["learning_rate_config"] = schedule
        values = []
        for step in steps:
            params = fresh_params(cfg)
            tf.GLOBAL_STEP[0] = np.asarray(step, np.int64)
            values.append(float(np.asarray(get_learning_rate(params, [], []).learning_rate)))
        lr_out[name] = {"schedule": schedule, "steps": steps, "values": values}
    with open(os.path.join(OUT, "reference_learning_rate.json"), "w") as f:
        json.dump(lr_out, f, indent=1)
    print("learning rate ok:", {k: v["values"][:3] for k, v in lr_out.items()})

    # ---- src/dataclass.py:ModelParameter on the shipped configs: everything it derives ---------------------------
    derived = {}
    configs = {name[:-5]: json.load(open(os.path.join(REFERENCE, "configs", name)))
               for name in sorted(os.listdir(os.path.join(REFERENCE, "configs"))) if name.endswith(".json")}
    configs["small"] = small_config()
    for name, config in configs.items():
        config = dict(config)
        if config.get("model_mode") == "gpt":  # src/main.py:87-93
            config.update(use_video=False, use_language=True)
        try:
            derived[name] = describe_parameters(fresh_params(config))
        except Exception as exc:  # noqa: BLE001
            print("dataclass", name, "FAILED:", type(exc).__name__, exc)
    with open(os.path.join(OUT, "reference_dataclass.json"), "w") as f:
        json.dump(derived, f, indent=1, sort_keys=True)
    print("dataclass ok:", sorted(derived))


if __name__ == "__main__":
    main()
