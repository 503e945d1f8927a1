"""Config / flag system of the hot path: the JSON surface of src/dataclass.py in the reference.

`ModelParameter(config_dict)` accepts the reference's model-config JSON (configs/*.json) unchanged: same keys,
same defaults, same derived sizes (features, intermediate, mesh/layout strings, named dimensions), same
warnings for unknown keys and the same ValueErrors (src/dataclass.py:184-241).  mesh_tensorflow types are
replaced by `utils_mtf.Dimension`; TensorFlow dtypes by torch dtypes.

Extensions (no counterpart in the reference, all optional):
  micro_batch_size  rows of the per-GPU batch processed per pass (the math is per-token; gradients are
                    accumulated in fp32), None = whole local batch
  gemm_backend      0 auto | 1 SIMT | 2 tcgen05   (include/obst_b200.h)
  device            torch device string
  fold_mid_norm     fold the group norm between W_mid and W_out into those two GEMMs (bf16, tcgen05 path): same
                    function, the normalised tensor is never written (model/basic.py).  Default true.
  keep_logits       a training step's `token_out` holds the logits (a copy: the loss kernel writes the logit gradient
                    over them in place); default false: `token_out` of a training step is that gradient buffer
  fold_mid_norm_backward  in the reversible backward also fold that norm's own backward into the GEMMs (the
                    gradient of W_out's input is never written either).  Default true.
  fuse_lm_head_loss  softmax cross-entropy, z-loss and accuracy taken in the LM head GEMM's epilogue when the
                    vocabulary is 256 wide (bf16, tcgen05 path, training): the logits are never written.  Default true.
  revnet_store_activations  keep every sub-block's intermediates instead of recomputing / reconstructing them in
                    backward (memory for speed; numerics differ from the reversible path only by the reference's
                    reconstruction rounding).  Default false = the reference's algorithm.
"""
import copy
import typing

import torch

from .utils_mtf import Dimension, Tensor

DTYPES = {"float32": torch.float32, "bfloat16": torch.bfloat16}


DEFAULTS: typing.Dict[str, typing.Any] = dict(
    bit_fold_value=4,
    block_config=[{'layer': ['norm-group-shift-scale', 'feed_forward-in_relu-group-in_glu_add-in_norm']}, {'layer': ['norm-group-std-shift-scale', 'attention-in_relu-embedded-relative']}],
    buffer_size=4, calc_accuracy=False, calculation_dtype='float32', color_channels=3, color_quantization_value=256,
    combine_assignments=False, concat_token=4, contrastive_across_samples=False,
    contrastive_across_token_embeddings=False, convolution_size=16, current_step=0, d_assignment=None,
    data_seed=456772, dataset_configs=[], debug_gradients=False, debug_sample=False, debug_train_step=False,
    default_sleep_duration=0.1, depth=16, embedding_stddev=0.04, empty_frame_embedding='absolute',
    equal_debugging_items_per_check=16, experts=64, features=None, features_per_head=None, frame_height=176,
    frame_width=320, grad_accumulation=1, group_linear_factor=2, heads=8, initial_autoregressive_position=128,
    input_block_config=[], input_dropout=0.0, interleaved_datasets=256, intermediate_feed_forward_multiplier=None,
    intermediate_feed_forward_multiplier_multiplier=None, iterations=2500, language_token_per_frame=0,
    learning_rate=5e-05, learning_rate_config={}, learning_rate_decay_min=5e-10, learning_rate_decay_multi=1,
    learning_rate_decay_start_step=100000, log_dict_keys=[], lookahead_alpha=0, lookahead_steps=0,
    macro_batch_loss_smoothing=False, macro_batching=1, masked_attention_dimensions=[0], max_checkpoints_keep=1,
    memory_reduction_strategy='revnet', mesh=None, mesh_impl=None, model_mode='jannet',
    model_path='gs://text-datasets/video-transformer/ctx=32-layer=64-heads=8-feat=256', momentum=0.95,
    momentumnet_alpha=0.99, multi_loss_strategy='linear', num_cores=0, num_cores_per_host=0, num_hosts=0,
    num_of_sample=10, opt_beta1=0.9, opt_beta2=0.999, optimizer='learning_rate',
    optimizer_calculation_dtype='float32', optimizer_slice_dtype='float32', other_color='\x1b[0m',
    output_block_config=[], output_embedding='absolute-orthogonal', output_offset=1, own_color='\x1b[32;1m',
    padding_token=0, parallel_batch=None, parallel_interleave=None, patch_size=16, pkm_axes=2,
    position_embedding='absolute', prefix='datasets/full_hd_video', reduce_lr_on_plateau_reduction=2,
    reduce_lr_on_plateau_timespan=0, rezero_lr_multiplier=0.1, sampling_temperature=0, save_graph=False,
    scale_by_depth=True, sequence_length=32, shuffle_buffer=256, shuffle_input_filenames=True,
    slice_dtype='float32', split_grad_accumulation=True, steps_per_checkpoint=100000, storage_dtype='float32',
    tensorflow_optimization_settings={}, three_axes=True, time_patch=1, token_embedding='absolute',
    token_patch_size=1, tpu_size=32, train=True, train_batch_size=1, train_steps=1073741824,
    use_autoregressive_sampling=False, use_bit_fold_input_pipeline=False, use_checkpointing=False,
    use_initial_position_embedding=False, use_language=True, use_random_dataloader=False, use_video=True,
    vocab_size=256, vocab_weight_factorization=0.125, warmup_steps=3000, web_workers=1, weight_centralisation=True,
    weight_decay=0.001, weight_standardisation=True, z_loss=0.0001,
)
# extensions of this implementation (see module docstring)
EXTENSIONS: typing.Dict[str, typing.Any] = dict(micro_batch_size=None, gemm_backend=0, device="cuda", init_seed=0,
                                                revnet_store_activations=False, fold_mid_norm=True,
                                                fold_mid_norm_backward=True, keep_logits=False,
                                                fuse_lm_head_loss=True)


class BlockConfig:
    def __init__(self, config, memory_reduction_strategy: str):
        if isinstance(config, BlockConfig):
            config = config.__dict__
        self.layer = []
        self.skip = False
        self.memory_reduction_strategy = memory_reduction_strategy
        self.__dict__.update(config)


class LearningRateConfig:
    def __init__(self, start_step: int = 0, final_step: int = 0, factor: float = 1.):
        self.start_step = start_step
        self.final_step = final_step
        self.factor = factor


class VariableDType(typing.NamedTuple):
    """(master/storage, slice, activation) triple of mtf.VariableDType."""
    master_dtype: torch.dtype
    slice_dtype: torch.dtype
    activation_dtype: torch.dtype


class TensorStorage:
    def __init__(self):
        self.text_input_embedding = None


_DTYPE_KEYS = ("slice_dtype", "storage_dtype", "calculation_dtype", "optimizer_slice_dtype", "optimizer_calculation_dtype")
_EMBEDDING_KEYS = ("position_embedding", "token_embedding", "output_embedding", "empty_frame_embedding")


class ModelParameter:
    """The reference's config object for the language ("gpt") path: every JSON key becomes an attribute, unknown keys
    warn, then the sizes and named dimensions the layers use are derived (src/dataclass.py:184-342).  What the
    reference derives only for its video / mixture-of-experts / TPU-mesh branches is left out: those branches are
    outside the hot path and `build` refuses them."""

    def __init__(self, config: typing.Dict[str, typing.Any]):
        values = copy.deepcopy(DEFAULTS)
        values.update(copy.deepcopy(EXTENSIONS))
        source = config.dict() if hasattr(config, "dict") and not isinstance(config, dict) else config
        for key, value in source.items():
            if key not in values:
                print(f"WARNING: Unknown ModelParameter {key}={value!r}")
            values[key] = value
        self.__dict__.update(values)
        self._validate()
        self._parse_strings()
        self._derive_sizes()
        self._name_dimensions()
        # ---- execution state of this implementation ----
        self.attention_idx = 0
        self.variable_cache = {}
        self.cached_parameters = {}
        self.tape: typing.Optional[list] = None     # backward closures recorded while not None
        self.residual = None                        # (buffer, sign, out) a final GEMM may fuse: revnet.py
        self.n_global_targets: typing.Optional[int] = None  # loss normaliser across DP ranks and micro-batches
        self.global_step = 0                        # tf.train global_step (int64 in the reference)
        self.manual_global_step = 0

    def _validate(self):
        """The ValueErrors of src/dataclass.py:189-241 that can fire on the language path."""
        if self.grad_accumulation > 1:
            raise ValueError("Gradient accumulation is not supported right now. The optimizer was split into two "
                             "different 'sections' where the 'accumulation' section still has to be integrated")
        if self.macro_batching < 1:
            raise ValueError("MacroBatching has to be >=1, where 1 means it's disabled")
        if not self.use_language and not self.use_video:
            raise ValueError("Language and video mode are disabled. No model can be built.")
        if self.features is None and self.features_per_head is None:
            raise ValueError("Either features or features_per_head has to be specified")

    def _parse_strings(self):
        for key in _EMBEDDING_KEYS:  # "absolute-orthogonal" -> ["absolute", "orthogonal"]
            value = getattr(self, key)
            if isinstance(value, str):
                setattr(self, key, value.split('-'))
        for key in _DTYPE_KEYS:
            name = getattr(self, key)
            if isinstance(name, torch.dtype):
                continue
            if name not in DTYPES:
                raise ValueError(f"{key}={name!r}: supported dtypes on this path are {list(DTYPES)}")
            setattr(self, key, DTYPES[name])
        self.learning_rate_config = {key: cfg if isinstance(cfg, LearningRateConfig) else LearningRateConfig(**cfg)
                                     for key, cfg in self.learning_rate_config.items()}
        self.variable_dtype = VariableDType(self.storage_dtype, self.slice_dtype, self.calculation_dtype)
        self.optimizer_dtype = VariableDType(self.storage_dtype, self.optimizer_slice_dtype,
                                             self.optimizer_calculation_dtype)
        self.tensor_storage = TensorStorage()
        strategy = self.memory_reduction_strategy
        self.block_config = [BlockConfig(c, memory_reduction_strategy=strategy) for c in self.block_config]
        self.input_block_config = [BlockConfig(c, memory_reduction_strategy="checkpoint")
                                   for c in self.input_block_config]
        self.output_block_config = [BlockConfig(c, memory_reduction_strategy="checkpoint")
                                    for c in self.output_block_config]

    def _derive_sizes(self):
        if self.features is None:
            self.features = self.features_per_head * self.heads
        if self.features_per_head is None:
            self.features_per_head = self.features // self.heads
        # width of the bottleneck: features * group_linear_factor * mm / heads  (src/dataclass.py:230-235,283-285)
        mm = self.intermediate_feed_forward_multiplier_multiplier
        if mm is not None:
            self.intermediate_feed_forward_multiplier = self.group_linear_factor * mm / self.heads
        elif self.intermediate_feed_forward_multiplier is None:
            self.intermediate_feed_forward_multiplier = self.group_linear_factor / self.heads
        if not self.use_video and self.language_token_per_frame != self.sequence_length:
            print(f"language_token_per_frame is unused in language-only mode. Overwriting with "
                  f"sequence_length={self.sequence_length}")
            self.language_token_per_frame = self.sequence_length
        self.time_patch_size = self.sequence_length // self.time_patch
        self.language_token_patch = self.language_token_per_frame // self.token_patch_size
        # The reference turns (tpu_size, heads) into an MTF mesh "b:<tpu_size//heads>,h:<heads>" with layout
        # "batch:b,heads:h" (src/dataclass.py:247-252).  The strings are kept for inspection only: here `batch:b` is
        # data parallelism over the GPUs of one box and all heads stay on one GPU.
        axes = [("b", self.tpu_size // self.heads, "batch")] * (self.heads < self.tpu_size) + \
               [("h", self.heads, "heads")] * (self.heads > 1)
        self.mesh_shape = ','.join(f"{a}:{n:.0f}" for a, n, _ in axes)
        self.layout = ','.join(f"{d}:{a}" for a, _, d in axes)

    def _name_dimensions(self):
        self.head_dim = Dimension("heads", self.heads)
        self.head_dimensions = [self.head_dim]
        self.key_dim = Dimension("features_per_head", self.features // self.heads)
        self.feature_dims = self.head_dimensions + [self.key_dim]
        width = int(self.heads * self.key_dim.size * self.intermediate_feed_forward_multiplier)
        self.intermediate = [Dimension("intermediate", width)]
        self.vocab_dim = Dimension('vocab', self.vocab_size)
        self.batch_dim = Dimension("batch", self.train_batch_size)
        self.macro_batch_dim = Dimension("batch", self.train_batch_size * self.macro_batching)
        self.sequence_dim = Dimension("sequence", self.time_patch_size)
        self.token_patch_dim = Dimension("language_token_patch", self.token_patch_size)
        self.token_dim_shape = [self.batch_dim, self.sequence_dim, self.token_patch_dim]
        self.input_pipeline_shape = {}
        if self.use_language:
            self.input_pipeline_shape['token_x'] = self.token_dim_shape
            self.input_pipeline_shape['token_y'] = self.token_dim_shape

    def dict(self) -> typing.Dict[str, typing.Any]:
        return self.__dict__

    def items(self):
        return self.__dict__.items()

    def values(self):
        return self.__dict__.values()

    def __str__(self) -> str:
        return str(self.__dict__)

    __repr__ = __str__


def gpt_mode(params_or_config):
    """The `model_mode == 'gpt'` override of src/main.py:87-93: language only."""
    cfg = dict(params_or_config)
    if cfg.get("model_mode", "jannet") == "gpt":
        cfg["use_video"] = False
        cfg["use_language"] = True
    return cfg


class BlockArgs:
    """Layer call context (src/dataclass.py:387-419).  NOTE the two reference quirks that are kept on purpose:
    `__call__` rebuilds the object WITHOUT `is_last` (so OrthogonalInit.scale_by_depth never fires on this path),
    and a `str` argument appends the *type* `str` rather than the string."""

    def __init__(self, params: ModelParameter, tensor: typing.Optional[Tensor], name_extras: typing.List[str],
                 is_last: bool = False):
        self.params = params
        self.tensor = tensor
        self.name_extras = name_extras
        self.is_last = is_last

    def __call__(self, *args):
        new = BlockArgs(self.params, self.tensor, self.name_extras[:])
        for a in args:
            if isinstance(a, ModelParameter):
                new.params = a
            elif isinstance(a, Tensor):
                new.tensor = a
            elif isinstance(a, (list, tuple)):
                new.name_extras = list(a)
            elif isinstance(a, str):
                new.name_extras.append(str)
            else:
                raise ValueError(f"Argument {a} is of unsupported type {type(a)}. "
                                 f"Only ModelParameter, Tensor, typing.List[str] and str are supported")
        return new

    def __iter__(self):
        for itm in self.name_extras:
            yield itm

    def __len__(self):
        return len(self.name_extras)

    def __getitem__(self, idx: int):
        return self.name_extras[idx]
