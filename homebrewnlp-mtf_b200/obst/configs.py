"""The three benchmark shapes of the reference (configs/32big_mixer.json, 32ctx_mixer.json, 32mixer_group.json:
identical except sequence_length / features_per_head / train_batch_size) plus BASELINE.json's tiny CPU-runnable
case, as config dicts.  Any reference JSON loaded with json.load works the same way (`ModelParameter(cfg)`)."""
import copy
import typing

_MIXER_BLOCKS = [
    {"layer": ["norm-shift-scale-features-group",
               "bottleneck_group_linear-in:relu-mid:relu-mid:norm-mid:shift-mid:scale-mid:features"]},
    {"layer": ["norm-shift-scale-features-group",
               "attention-biased_attention_map-absolute-input_as_value-shared",
               "norm-shift-scale-features-group",
               "activation-gelu",
               "attention-biased_attention_map-absolute-input_as_value-shared"]},
]

_SHAPES = {  # name: (sequence_length, features_per_head, train_batch_size, depth, dtype)
    "32big_mixer": (512, 512, 1024, 32, "bfloat16"),
    "32ctx_mixer": (2048, 256, 256, 32, "bfloat16"),
    "32mixer_group": (256, 256, 4096, 32, "bfloat16"),
    "tiny": (256, 32, 8, 2, "float32"),
}


def mixer_config(name: str = "32big_mixer", **overrides) -> typing.Dict[str, typing.Any]:
    seq, fph, batch, depth, dtype = _SHAPES[name]
    cfg = {
        "model_mode": "gpt", "use_language": True, "use_video": False,
        "embedding_stddev": 0.004, "calc_accuracy": True, "scale_by_depth": True,
        "block_config": copy.deepcopy(_MIXER_BLOCKS),
        "group_linear_factor": 2, "intermediate_feed_forward_multiplier_multiplier": 0.5,
        "depth": depth, "use_initial_position_embedding": False,
        "sequence_length": seq, "features_per_head": fph, "heads": 8, "train_batch_size": batch,
        "data_seed": 134567, "vocab_size": 256,
        "learning_rate": 0.01, "opt_beta1": 0.9, "opt_beta2": 0.99, "memory_reduction_strategy": "revnet",
        "optimizer": "adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", "weight_decay": 0.0001,
        "weight_centralisation": False, "weight_standardisation": False,
        "macro_batching": 1, "macro_batch_loss_smoothing": True, "use_checkpointing": False,
        "calculation_dtype": dtype, "storage_dtype": dtype, "optimizer_slice_dtype": dtype,
        "slice_dtype": "float32",
        "learning_rate_config": {"linear_warmup": {"final_step": 4096}},
    }
    cfg.update(overrides)
    return cfg
