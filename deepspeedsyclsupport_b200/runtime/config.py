"""`DeepSpeedConfig`: parse ds_config (path to JSON/HJSON, dict, or base64-encoded JSON), validate it and solve
the batch-size triad.  Parity: deepspeed/runtime/config.py:696-1027 (_initialize_params :789-900,
_set_batch_related_parameters :919-962, _batch_assertion :902-917, elastic override :722-782)."""
import base64
import copy
import json
import os
from typing import Optional

import torch

from ..comm.config import DeepSpeedCommsConfig
from ..utils.logging import logger
from . import constants as C
from .config_utils import (DeepSpeedConfigModel, ScientificNotationEncoder, dict_raise_error_on_duplicate_keys,
                           get_scalar_param)
from .zero.config import ZERO_OPTIMIZATION, ZeroStageEnum, get_zero_config

ADAGRAD_OPTIMIZER = "adagrad"
ADAM_OPTIMIZER = "adam"
ADAMW_OPTIMIZER = "adamw"
LAMB_OPTIMIZER = "lamb"
ONEBIT_ADAM_OPTIMIZER = "onebitadam"
ZERO_ONE_ADAM_OPTIMIZER = "zerooneadam"
ONEBIT_LAMB_OPTIMIZER = "onebitlamb"
MUADAM_OPTIMIZER = "muadam"
MUADAMW_OPTIMIZER = "muadamw"
MUSGD_OPTIMIZER = "musgd"
LION_OPTIMIZER = "lion"
SGD_OPTIMIZER = "sgd"
DEEPSPEED_OPTIMIZERS = [ADAGRAD_OPTIMIZER, ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER,
                        ONEBIT_LAMB_OPTIMIZER, ZERO_ONE_ADAM_OPTIMIZER, MUADAM_OPTIMIZER, MUADAMW_OPTIMIZER,
                        MUSGD_OPTIMIZER, LION_OPTIMIZER]
TORCH_ADAM_PARAM = "torch_adam"
ADAM_W_MODE = "adam_w_mode"
ADAM_W_MODE_DEFAULT = True


class DeepSpeedConfigError(Exception):
    pass


class DtypeEnum:
    _map = {"fp16": torch.float16, "float16": torch.float16, "half": torch.float16, "fp32": torch.float32,
            "float32": torch.float32, "float": torch.float32, "bf16": torch.bfloat16, "bfloat16": torch.bfloat16,
            "int8": torch.int8}

    @classmethod
    def get(cls, name):
        if name is None or isinstance(name, torch.dtype):
            return name
        if name not in cls._map:
            raise ValueError(f"Invalid dtype '{name}'; supported: {sorted(cls._map)}")
        return cls._map[name]


# ----------------------------------------------------------------------------- small sub-blocks
class ActivationCheckpointingConfig(DeepSpeedConfigModel):
    """parity: runtime/activation_checkpointing/config.py"""
    partition_activations: bool = False
    contiguous_memory_optimization: bool = False
    cpu_checkpointing: bool = False
    number_checkpoints: Optional[int] = None
    synchronize_checkpoint_boundary: bool = False
    profile: bool = False


class AioConfig(DeepSpeedConfigModel):
    """parity: runtime/swap_tensor/aio_config.py"""
    block_size: int = 1048576
    queue_depth: int = 8
    thread_count: int = 1
    single_submit: bool = False
    overlap_events: bool = True


class FlopsProfilerConfig(DeepSpeedConfigModel):
    """parity: profiling/config.py"""
    enabled: bool = False
    recompute_fwd_factor: float = 0.0
    profile_step: int = 1
    module_depth: int = -1
    top_modules: int = 1
    detailed: bool = True
    output_file: Optional[str] = None


class TensorBoardConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class WandbConfig(DeepSpeedConfigModel):
    enabled: bool = False
    group: Optional[str] = None
    team: Optional[str] = None
    project: str = "deepspeed"


class CSVConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class DeepSpeedMonitorConfig(DeepSpeedConfigModel):
    """parity: monitor/config.py:15-79"""
    tensorboard: TensorBoardConfig = TensorBoardConfig()
    wandb: WandbConfig = WandbConfig()
    csv_monitor: CSVConfig = CSVConfig()

    @property
    def enabled(self):
        return self.tensorboard.enabled or self.wandb.enabled or self.csv_monitor.enabled


class PipelineConfig(DeepSpeedConfigModel):
    stages: str = "auto"
    partition: str = "best"
    seed_layers: bool = False
    activation_checkpoint_interval: int = 0
    pipe_partitioned: bool = True
    grad_partitioned: bool = True
    use_reentrant: bool = True


class HybridEngineConfig(DeepSpeedConfigModel):
    enabled: bool = False
    max_out_tokens: int = 512
    inference_tp_size: int = 1
    release_inference_cache: bool = False
    pin_parameters: bool = True
    tp_gather_partition_size: int = 8


class NebulaConfig(DeepSpeedConfigModel):
    enabled: bool = False
    persistent_storage_path: Optional[str] = None
    persistent_time_interval: int = 100
    num_of_version_in_retention: int = 2
    enable_nebula_load: bool = True
    load_path: Optional[str] = None


class DeepSpeedConfigWriter:

    def __init__(self, data=None):
        self.data = data if data is not None else {}

    def add_config(self, key, value):
        self.data[key] = value

    def load_config(self, filename):
        self.data = json.load(open(filename, "r"), object_pairs_hook=dict_raise_error_on_duplicate_keys)

    def write_config(self, filename):
        with open(filename, "w") as outfile:
            json.dump(self.data, outfile)


def _load_config_dict(config):
    if isinstance(config, dict):
        return copy.deepcopy(config)
    if isinstance(config, (str, os.PathLike)) and os.path.exists(str(config)):
        text = open(str(config), "r").read()
        try:
            return json.loads(text, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        except json.JSONDecodeError:
            import hjson
            return hjson.loads(text, object_pairs_hook=dict_raise_error_on_duplicate_keys)
    if isinstance(config, str):
        try:
            config_decoded = base64.urlsafe_b64decode(config).decode("utf-8")
            return json.loads(config_decoded, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        except (UnicodeDecodeError, AttributeError, ValueError, base64.binascii.Error):
            raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary or a valid "
                             f"base64. Received: {config}")
    raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary. Received: {config}")


class DeepSpeedConfig:

    def __init__(self, config, mpu=None, world_size=None):
        self._param_dict = _load_config_dict(config)
        try:
            from .. import comm as dist
            self.global_rank = dist.get_rank()
            if world_size is not None:
                self.world_size = world_size
            elif mpu is None:
                self.world_size = dist.get_world_size()
            else:
                self.world_size = mpu.get_data_parallel_world_size()
        except Exception:
            self.global_rank = 0
            self.world_size = world_size or 1
        # sequence parallel ranks share a batch: dp world excludes them (engine passes the right mpu)
        self.elasticity_enabled = False
        self._elastic_override()
        self._initialize_params(copy.copy(self._param_dict))
        self._configure_train_batch_size()
        self._do_sanity_check()

    # -- elasticity: override the batch triad with the pre-computed elastic batch (config.py:722-782)
    def _elastic_override(self):
        from ..elasticity import (elasticity_enabled, compute_elastic_config, ensure_immutable_elastic_config)
        from ..elasticity.constants import ELASTICITY, IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT
        from ..version import __version__
        if not elasticity_enabled(self._param_dict):
            return
        self.elasticity_enabled = True
        logger.info("DeepSpeed elasticity support enabled")
        final_batch_size, valid_gpus, micro_batch_size = compute_elastic_config(
            ds_config=self._param_dict, target_deepspeed_version=__version__, world_size=self.world_size)
        elastic_dict = self._param_dict[ELASTICITY]
        ensure_immutable_elastic_config(runtime_elastic_config_dict=elastic_dict)
        ignore = elastic_dict.get(IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT)
        if not ignore:
            for k in (C.TRAIN_BATCH_SIZE, C.TRAIN_MICRO_BATCH_SIZE_PER_GPU, C.GRADIENT_ACCUMULATION_STEPS):
                if k in self._param_dict:
                    raise DeepSpeedConfigError(
                        f"One or more batch related parameters were found in your ds_config ({k}); they are "
                        "overridden by elasticity. Set 'ignore_non_elastic_batch_info': true to silence.")
        gas = final_batch_size // (micro_batch_size * self.world_size)
        self._param_dict[C.TRAIN_BATCH_SIZE] = final_batch_size
        self._param_dict[C.TRAIN_MICRO_BATCH_SIZE_PER_GPU] = micro_batch_size
        self._param_dict[C.GRADIENT_ACCUMULATION_STEPS] = gas

    def _initialize_params(self, pd):
        g = get_scalar_param
        self.train_batch_size = g(pd, C.TRAIN_BATCH_SIZE, None)
        self.train_micro_batch_size_per_gpu = g(pd, C.TRAIN_MICRO_BATCH_SIZE_PER_GPU, None)
        self.gradient_accumulation_steps = g(pd, C.GRADIENT_ACCUMULATION_STEPS, None)
        self.steps_per_print = g(pd, C.STEPS_PER_PRINT, C.STEPS_PER_PRINT_DEFAULT)
        self.dump_state = g(pd, C.DUMP_STATE, C.DUMP_STATE_DEFAULT)
        self.disable_allgather = g(pd, C.DISABLE_ALLGATHER, C.DISABLE_ALLGATHER_DEFAULT)
        self.communication_data_type = DtypeEnum.get(g(pd, C.COMMUNICATION_DATA_TYPE, None))
        self.seq_parallel_communication_data_type = DtypeEnum.get(
            g(pd, C.SEQ_PARALLEL_COMMUNICATION_DATA_TYPE, C.SEQ_PARALLEL_COMMUNICATION_DATA_TYPE_DEFAULT))
        self.prescale_gradients = g(pd, C.PRESCALE_GRADIENTS, C.PRESCALE_GRADIENTS_DEFAULT)
        self.gradient_predivide_factor = g(pd, C.GRADIENT_PREDIVIDE_FACTOR, C.GRADIENT_PREDIVIDE_FACTOR_DEFAULT)
        self.sparse_gradients_enabled = g(pd, C.SPARSE_GRADIENTS, C.SPARSE_GRADIENTS_DEFAULT)

        self.zero_config = get_zero_config(pd)
        self.mics_shard_size = self.zero_config.mics_shard_size
        self.mics_hierarchial_params_gather = self.zero_config.mics_hierarchical_params_gather
        self.zero_optimization_stage = int(self.zero_config.stage)
        self.zero_enabled = self.zero_optimization_stage > 0

        self.activation_checkpointing_config = ActivationCheckpointingConfig(**pd.get("activation_checkpointing", {}))
        self.comms_config = DeepSpeedCommsConfig(pd)
        self.monitor_config = DeepSpeedMonitorConfig(**{k: pd[k] for k in (C.TENSORBOARD, C.WANDB, C.CSV_MONITOR) if k in pd})

        self.gradient_clipping = g(pd, C.GRADIENT_CLIPPING, C.GRADIENT_CLIPPING_DEFAULT)
        fp16 = pd.get(C.FP16, {})
        self.fp16_enabled = fp16.get(C.FP16_ENABLED, False)
        self.fp16_auto_cast = fp16.get(C.FP16_AUTO_CAST, False) if self.fp16_enabled else None
        bf = pd.get(C.BFLOAT16, pd.get(C.BFLOAT16_OLD, {}))
        self.bfloat16_enabled = bf.get(C.BFLOAT16_ENABLED, False)
        self.bfloat16_immediate_grad_update = bf.get(C.BFLOAT16_IMMEDIATE_GRAD_UPDATE, False)
        assert not (self.fp16_enabled and self.bfloat16_enabled), "bfloat16 and fp16 modes cannot be simultaneously enabled"
        self.fp16_master_weights_and_gradients = fp16.get(C.FP16_MASTER_WEIGHTS_AND_GRADS, False) if self.fp16_enabled else False
        self.amp_enabled = pd.get(C.AMP, {}).get(C.AMP_ENABLED, False)
        self.amp_params = {k: v for k, v in pd.get(C.AMP, {}).items() if k != C.AMP_ENABLED} if self.amp_enabled else False
        self.loss_scale = fp16.get(C.FP16_LOSS_SCALE, C.FP16_LOSS_SCALE_DEFAULT) if self.fp16_enabled else 1.0
        self.initial_dynamic_scale = 2 ** fp16.get(C.FP16_INITIAL_SCALE_POWER, C.FP16_INITIAL_SCALE_POWER_DEFAULT)
        self.dynamic_loss_scale_args = None
        if self.fp16_enabled and any(k in fp16 for k in (C.FP16_INITIAL_SCALE_POWER, C.FP16_LOSS_SCALE_WINDOW,
                                                          C.FP16_MIN_LOSS_SCALE, C.FP16_HYSTERESIS,
                                                          C.FP16_CONSECUTIVE_HYSTERESIS)):
            self.dynamic_loss_scale_args = {
                "init_scale": 2 ** fp16.get(C.FP16_INITIAL_SCALE_POWER, C.FP16_INITIAL_SCALE_POWER_DEFAULT),
                "scale_window": fp16.get(C.FP16_LOSS_SCALE_WINDOW, C.FP16_LOSS_SCALE_WINDOW_DEFAULT),
                "delayed_shift": fp16.get(C.FP16_HYSTERESIS, C.FP16_HYSTERESIS_DEFAULT),
                "consecutive_hysteresis": fp16.get(C.FP16_CONSECUTIVE_HYSTERESIS, C.FP16_CONSECUTIVE_HYSTERESIS_DEFAULT),
                "min_scale": fp16.get(C.FP16_MIN_LOSS_SCALE, C.FP16_MIN_LOSS_SCALE_DEFAULT),
            }

        from ..compression.config import get_compression_config, get_quantize_enabled
        self.compression_config = get_compression_config(pd)
        self.graph_harvesting = g(pd, C.GRAPH_HARVESTING, C.GRAPH_HARVESTING_DEFAULT)

        opt = pd.get(C.OPTIMIZER, {})
        self.optimizer_name = opt.get(C.TYPE, None)
        if self.optimizer_name is not None and self.optimizer_name.lower() in DEEPSPEED_OPTIMIZERS:
            self.optimizer_name = self.optimizer_name.lower()
        self.optimizer_params = opt.get(C.OPTIMIZER_PARAMS, None)
        self.optimizer_legacy_fusion = opt.get(C.LEGACY_FUSION, C.LEGACY_FUSION_DEFAULT)
        self.zero_allow_untested_optimizer = g(pd, C.ZERO_ALLOW_UNTESTED_OPTIMIZER, C.ZERO_ALLOW_UNTESTED_OPTIMIZER_DEFAULT)
        self.zero_force_ds_cpu_optimizer = g(pd, C.ZERO_FORCE_DS_CPU_OPTIMIZER, C.ZERO_FORCE_DS_CPU_OPTIMIZER_DEFAULT)
        sch = pd.get(C.SCHEDULER, {})
        self.scheduler_name = sch.get(C.TYPE, None)
        self.scheduler_params = sch.get(C.SCHEDULER_PARAMS, None)

        self.flops_profiler_config = FlopsProfilerConfig(**pd.get("flops_profiler", {}))
        self.wall_clock_breakdown = g(pd, C.WALL_CLOCK_BREAKDOWN, C.WALL_CLOCK_BREAKDOWN_DEFAULT) | self.flops_profiler_config.enabled
        self.memory_breakdown = g(pd, C.MEMORY_BREAKDOWN, C.MEMORY_BREAKDOWN_DEFAULT)
        from ..autotuning.config import DeepSpeedAutotuningConfig
        self.autotuning_config = DeepSpeedAutotuningConfig(**pd.get("autotuning", {}))

        self.sparse_attention = pd.get(C.SPARSE_ATTENTION, None)
        self.pipeline = PipelineConfig(**pd.get(C.PIPELINE, {})).model_dump()
        pld = pd.get(C.PROGRESSIVE_LAYER_DROP, {})
        self.pld_enabled = pld.get(C.PLD_ENABLED, False)
        self.pld_params = {k: v for k, v in pld.items() if k != C.PLD_ENABLED} if self.pld_enabled else False
        cl = pd.get(C.CURRICULUM_LEARNING_LEGACY, {})
        self.curriculum_enabled_legacy = cl.get("enabled", False)
        self.curriculum_params_legacy = {k: v for k, v in cl.items() if k != "enabled"} if self.curriculum_enabled_legacy else False
        from .data_pipeline.config import get_data_efficiency_config
        self.data_efficiency_enabled = pd.get(C.DATA_EFFICIENCY, {}).get("enabled", False)
        self.data_efficiency_config = get_data_efficiency_config(pd)

        ckpt = pd.get(C.CHECKPOINT, {})
        mode = str(ckpt.get(C.CHECKPOINT_TAG_VALIDATION, C.CHECKPOINT_TAG_VALIDATION_DEFAULT)).upper()
        valid = [m.upper() for m in C.CHECKPOINT_TAG_VALIDATION_MODES]
        if mode not in valid:
            raise DeepSpeedConfigError(f"Checkpoint config contains invalid tag_validation value of {mode}, expecting one of {valid}")
        self.checkpoint_tag_validation_enabled = mode != "IGNORE"
        self.checkpoint_tag_validation_fail = mode == "FAIL"
        self.load_universal_checkpoint = ckpt.get(C.LOAD_UNIVERSAL_CHECKPOINT, C.LOAD_UNIVERSAL_CHECKPOINT_DEFAULT)
        self.use_node_local_storage = ckpt.get(C.USE_NODE_LOCAL_STORAGE_CHECKPOINT, C.USE_NODE_LOCAL_STORAGE_CHECKPOINT_DEFAULT)
        self.checkpoint_parallel_write_pipeline = ckpt.get(C.CHECKPOINT_PARALLEL_WRITE, {}).get(
            C.CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE, C.CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE_DEFAULT)
        dt = pd.get(C.DATA_TYPES, {})
        self.grad_accum_dtype = dt.get(C.GRAD_ACCUM_DTYPE, C.GRAD_ACCUM_DTYPE_DEFAULT)

        self.aio_config = AioConfig(**pd.get("aio", {})).model_dump()
        self.dataloader_drop_last = g(pd, C.DATALOADER_DROP_LAST, C.DATALOADER_DROP_LAST_DEFAULT)
        self.nebula_config = NebulaConfig(**pd.get("nebula", {}))
        self.hybrid_engine = HybridEngineConfig(**pd.get("hybrid_engine", {}))
        self.weight_quantization_config = pd.get("weight_quantization", None)
        self.eigenvalue_enabled = pd.get(C.EIGENVALUE, {}).get("enabled", False)
        ev = pd.get(C.EIGENVALUE, {})
        self.eigenvalue_verbose = ev.get("verbose", False)
        self.eigenvalue_max_iter = ev.get("max_iter", 100)
        self.eigenvalue_tol = ev.get("tol", 1e-2)
        self.eigenvalue_stability = ev.get("stability", 1e-6)
        self.eigenvalue_gas_boundary_resolution = ev.get("gas_boundary_resolution", 1)
        self.eigenvalue_layer_name = ev.get("layer_name", "bert.encoder.layer")
        self.eigenvalue_layer_num = ev.get("layer_num", 0)
        self.use_data_before_expert_parallel_ = g(pd, C.USE_DATA_BEFORE_EXPERT_PARALLEL, C.USE_DATA_BEFORE_EXPERT_PARALLEL_DEFAULT)
        self.vocabulary_size = g(pd, C.VOCABULARY_SIZE, C.VOCABULARY_SIZE_DEFAULT)

    # -- batch triad: train_batch = micro_batch * gas * dp_world  (config.py:902-962)
    def _batch_assertion(self):
        tb, mb, gas = self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps
        assert tb > 0, f"Train batch size: {tb} has to be greater than 0"
        assert mb > 0, f"Micro batch size per gpu: {mb} has to be greater than 0"
        assert gas > 0, f"Gradient accumulation steps: {gas} has to be greater than 0"
        assert tb == mb * gas * self.world_size, (
            f"Check batch related parameters. train_batch_size is not equal to micro_batch_per_gpu * gradient_acc_step"
            f" * world_size {tb} != {mb} * {gas} * {self.world_size}")

    def _set_batch_related_parameters(self):
        tb, mb, gas = self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps
        ws = self.world_size
        if tb is not None and mb is not None and gas is not None:
            return
        if tb is not None and mb is not None:
            self.gradient_accumulation_steps = tb // mb // ws
        elif tb is not None and gas is not None:
            self.train_micro_batch_size_per_gpu = tb // ws // gas
        elif mb is not None and gas is not None:
            self.train_batch_size = mb * gas * ws
        elif tb is not None:
            self.gradient_accumulation_steps = 1
            self.train_micro_batch_size_per_gpu = tb // ws
        elif mb is not None:
            self.train_batch_size = mb * ws
            self.gradient_accumulation_steps = 1
        else:
            assert False, "Either train_batch_size or train_micro_batch_size_per_gpu needs to be provided"

    def _configure_train_batch_size(self):
        self._set_batch_related_parameters()
        self._batch_assertion()

    def _do_sanity_check(self):
        if self.zero_enabled:
            assert self.zero_optimization_stage <= int(ZeroStageEnum.max_stage), \
                f"DeepSpeedConfig: Maximum supported ZeRO stage is {int(ZeroStageEnum.max_stage)}"
        if self.fp16_master_weights_and_gradients:
            assert self.zero_enabled and self.zero_optimization_stage == 2, \
                "Fp16_master_weights_and_grads is only supported with ZeRO Stage 2 for now."
        if self.optimizer_params is not None and C.MAX_GRAD_NORM in self.optimizer_params:
            if self.fp16_enabled:
                logger.warning(f"DeepSpeedConfig: In FP16 mode, DeepSpeed will pass {C.MAX_GRAD_NORM} to FP16 wrapper")
            else:
                logger.warning(f"DeepSpeedConfig: In FP32 mode, {C.MAX_GRAD_NORM} is not permitted; setting to zero")
                self.optimizer_params[C.MAX_GRAD_NORM] = 0.0

    def print_user_config(self):
        logger.info("  json = {}".format(json.dumps(self._param_dict, sort_keys=True, indent=4,
                                                     cls=ScientificNotationEncoder, separators=(",", ":"))))

    def print(self, name):
        logger.info("{}:".format(name))
        for arg in sorted(vars(self)):
            if arg != "_param_dict":
                dots = "." * (29 - len(arg))
                logger.info("  {} {} {}".format(arg, dots, getattr(self, arg)))
        self.print_user_config()
