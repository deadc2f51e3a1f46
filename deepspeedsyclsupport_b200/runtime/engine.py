"""DeepSpeedEngine: the training wrapper returned by `initialize()`.

API parity target: deepspeed/runtime/engine.py:179-3600 (forward :1781, backward :1922, step :2120,
_take_model_step :2054, save_checkpoint :3050, load_checkpoint :2688, _zero3_consolidated_16bit_state_dict :3477,
config getters :473-879).  Internals are B200-first: every precision/ZeRO combination runs on the flat unit store
(`runtime/zero/sharded_optimizer.py`), there is no per-parameter bucketed all-reduce fallback, norms/clip/overflow are
computed on the device, and communication overlaps on dedicated CUDA streams.
"""
import hashlib
import os
import re
from collections import OrderedDict

import torch
from torch.nn.modules import Module
from torch.optim import Optimizer
from torch.optim.lr_scheduler import _LRScheduler

from .. import comm as dist
from ..accelerator import get_accelerator
from ..utils import groups
from ..utils.logging import log_dist, logger
from ..utils.timer import (BACKWARD_GLOBAL_TIMER, BACKWARD_INNER_GLOBAL_TIMER, BACKWARD_INNER_MICRO_TIMER,
                           BACKWARD_MICRO_TIMER, BACKWARD_REDUCE_GLOBAL_TIMER, BACKWARD_REDUCE_MICRO_TIMER,
                           FORWARD_GLOBAL_TIMER, FORWARD_MICRO_TIMER, STEP_GLOBAL_TIMER, STEP_MICRO_TIMER, NoopTimer,
                           SynchronizedWallClockTimer, ThroughputTimer)
from ..version import __version__
from . import lr_schedules
from .config import (ADAGRAD_OPTIMIZER, ADAM_OPTIMIZER, ADAMW_OPTIMIZER, ADAM_W_MODE, ADAM_W_MODE_DEFAULT, DEEPSPEED_OPTIMIZERS,
                     LAMB_OPTIMIZER, LION_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER, ONEBIT_LAMB_OPTIMIZER, SGD_OPTIMIZER,
                     TORCH_ADAM_PARAM, ZERO_ONE_ADAM_OPTIMIZER, DeepSpeedConfig)
from .dataloader import DeepSpeedDataLoader
from .zero.config import ZeroStageEnum
from .zero.partition_parameters import GatheredParameters
from .zero.param_store import ZeroParamStatus

MEMORY_OPT_ALLREDUCE_SIZE = 500000000
DeepSpeedOptimizerCallable = None
DeepSpeedSchedulerCallable = None


from .engine_compat import EngineCompatMixin

class EngineTimers:
    """Names of the wall-clock timers the engine maintains (parity: engine.py:144-176)."""

    def __init__(self, enable_micro_timers, enable_global_timers):
        self.forward_timers = []
        self.backward_timers = []
        self.backward_inner_timers = []
        self.backward_reduce_timers = []
        self.step_timers = []
        self.global_timers = []
        self.micro_timers = []
        if enable_micro_timers:
            self.forward_timers += [FORWARD_MICRO_TIMER]
            self.backward_timers += [BACKWARD_MICRO_TIMER]
            self.backward_inner_timers += [BACKWARD_INNER_MICRO_TIMER]
            self.backward_reduce_timers += [BACKWARD_REDUCE_MICRO_TIMER]
            self.step_timers += [STEP_MICRO_TIMER]
            self.micro_timers += [FORWARD_MICRO_TIMER, BACKWARD_MICRO_TIMER, BACKWARD_INNER_MICRO_TIMER,
                                  BACKWARD_REDUCE_MICRO_TIMER, STEP_MICRO_TIMER]
        if enable_global_timers:
            self.forward_timers += [FORWARD_GLOBAL_TIMER]
            self.backward_timers += [BACKWARD_GLOBAL_TIMER]
            self.backward_inner_timers += [BACKWARD_INNER_GLOBAL_TIMER]
            self.backward_reduce_timers += [BACKWARD_REDUCE_GLOBAL_TIMER]
            self.step_timers += [STEP_GLOBAL_TIMER]
            self.global_timers += [FORWARD_GLOBAL_TIMER, BACKWARD_GLOBAL_TIMER, BACKWARD_INNER_GLOBAL_TIMER,
                                   BACKWARD_REDUCE_GLOBAL_TIMER, STEP_GLOBAL_TIMER]


class DeepSpeedEngine(EngineCompatMixin, Module):

    def __init__(self, args, model, optimizer=None, model_parameters=None, training_data=None, lr_scheduler=None, mpu=None,
                 dist_init_required=None, collate_fn=None, config=None, config_class=None, dont_change_device=False):
        super().__init__()
        self.dont_change_device = dont_change_device
        self.client_optimizer = optimizer
        self.client_lr_scheduler = lr_scheduler
        self.training_data = training_data
        self.collate_fn = collate_fn
        self.mpu = mpu
        self.all_to_all_group = None
        self.data_parallel_group = None
        self.global_steps = 0
        self.global_samples = 0
        self.micro_steps = 0
        self.skipped_steps = 0
        self.gradient_average = True
        self.warn_unscaled_loss = True
        self.config = config
        self._config = config_class
        self.loaded_checkpoint_mp_world_size = None
        self.loaded_checkpoint_dp_world_size = None
        self.enable_backward_allreduce = True
        self.progressive_layer_drop = None
        self.eigenvalue = None
        self.block_eigenvalue = None
        self.gas_boundary_ctr = 0
        self.dist_backend = get_accelerator().communication_backend_name()
        self.has_moe_layers = False
        self.num_experts = []
        self.gate_modules = []
        self.moe_layers = []
        self._step_applied = False
        self._global_grad_norm = None
        self.use_ds_comm = False
        self.checkpoint_engine = None
        self._is_gradient_accumulation_boundary = None
        self.scale_wrt_gas = None
        self.losses = None
        self.optimizer = None
        self.basic_optimizer = None
        self.lr_scheduler = None
        self._is_compiled = False

        if dist_init_required is not False and not dist.is_initialized():
            dist.init_distributed(dist_backend=self.dist_backend)
        self._do_args_sanity_check(args)
        self._configure_with_arguments(args, mpu)
        self._do_sanity_check()
        self._set_distributed_vars(args)
        dist.configure(self._config)
        from ..monitor.monitor import MonitorMaster
        self.monitor = MonitorMaster(self._config.monitor_config)

        self.module = model
        self._configure_distributed_model(model)
        self.param_names = {param: name for name, param in model.named_parameters()}
        for name, p in model.named_parameters():
            p.ds_name = name

        self.wall_clock_breakdown_ = self._config.wall_clock_breakdown
        self.timers = SynchronizedWallClockTimer() if self.wall_clock_breakdown() else NoopTimer()
        self.engine_timers = EngineTimers(enable_micro_timers=self.wall_clock_breakdown(),
                                          enable_global_timers=self.wall_clock_breakdown() or self.flops_profiler_enabled())
        self.tput_timer = ThroughputTimer(batch_size=self.train_batch_size(), steps_per_output=self.steps_per_print(),
                                          monitor_memory=False)
        self.flops_profiler = None
        if self.flops_profiler_enabled():
            from ..profiling.flops_profiler.profiler import FlopsProfiler
            self.flops_profiler = FlopsProfiler(self.module, self, self.flops_profiler_recompute_fwd_factor())

        self.training_dataloader = self.deepspeed_io(training_data) if training_data is not None else None

        model_parameters = list(model_parameters) if model_parameters is not None else None
        has_optimizer = optimizer is not None or self.optimizer_name() is not None
        if has_optimizer:
            self._configure_optimizer(optimizer, model_parameters)
            self._configure_lr_scheduler(lr_scheduler)
        elif self.zero_optimization_stage() == 3 or self.bfloat16_enabled() or self.fp16_enabled():
            # inference / evaluation under ZeRO-3: parameters are sharded, no optimizer state is created
            self.optimizer = self._wrap_zero(None) if self.zero_optimization_stage() == 3 else None

        self._configure_checkpointing()
        if self.pld_enabled():
            from .progressive_layer_drop import ProgressiveLayerDrop
            self.progressive_layer_drop = ProgressiveLayerDrop(**self.pld_params())
        if self.curriculum_enabled_legacy():
            from .data_pipeline.curriculum_scheduler import CurriculumScheduler
            self.curriculum_scheduler_legacy = CurriculumScheduler(self.curriculum_params_legacy())
        self.random_ltd_scheduler = None
        self.training_data_sampler = None
        if self.random_ltd_enabled():
            from .data_pipeline import constants as K
            from .data_pipeline.data_routing.scheduler import RandomLTDScheduler
            ltd = self.random_ltd_config()
            ltd[K.RANDOM_LTD_GLOBAL_BATCH_SIZE] = self.train_batch_size()
            ltd[K.RANDOM_LTD_MICRO_BATCH_SIZE] = self.train_micro_batch_size_per_gpu()
            self.random_ltd_scheduler = RandomLTDScheduler(ltd)
        if self.eigenvalue_enabled():
            from .eigenvalue import Eigenvalue
            c = self._config
            self.eigenvalue = Eigenvalue(verbose=c.eigenvalue_verbose, max_iter=c.eigenvalue_max_iter, tol=c.eigenvalue_tol,
                                         stability=c.eigenvalue_stability, gas_boundary_resolution=c.eigenvalue_gas_boundary_resolution,
                                         layer_name=c.eigenvalue_layer_name, layer_num=c.eigenvalue_layer_num)
        self.compression_scheduler = None
        cc = self._config.compression_config
        if cc is not None and any(v.get("shared_parameters", {}).get("enabled", False) for k, v in cc.items() if k != "layer_reduction"):
            from ..compression.scheduler import compression_scheduler
            self.compression_scheduler = compression_scheduler(self.module, cc)
        self.quantizer = self._configure_quantization()
        if self.global_rank == 0 and self._config.dump_state:
            self._config.print("DeepSpeedEngine configuration")
        self.unflatten = torch._utils._unflatten_dense_tensors
        self.flatten = torch._utils._flatten_dense_tensors

    # ------------------------------------------------------------------ attribute passthrough
    def __getattr__(self, name):
        _module = {}
        if "module" in self.__dict__.get("_modules", {}):
            _module = self.__dict__["_modules"]["module"]
        try:
            return super().__getattr__(name)
        except AttributeError:
            if _module is not {} and hasattr(_module, name):
                return getattr(_module, name)
            raise AttributeError(f"'{type(self).__name__}' object has no attribute '{name}'")

    # ------------------------------------------------------------------ config getters (engine.py:473-879)
    def train_batch_size(self):
        return self._config.train_batch_size

    def train_micro_batch_size_per_gpu(self):
        return self._config.train_micro_batch_size_per_gpu

    def gradient_accumulation_steps(self):
        return self._config.gradient_accumulation_steps

    def set_train_batch_size(self, train_batch_size):
        if train_batch_size % (self.train_micro_batch_size_per_gpu() * self.dp_world_size) != 0:
            raise ValueError("Train batch size must be divisible by micro-batch data parallelism")
        new_gas = train_batch_size // (self.train_micro_batch_size_per_gpu() * self.dp_world_size)
        self._config.train_batch_size = train_batch_size
        self._config.gradient_accumulation_steps = new_gas

    def set_train_micro_batch_size(self, micro_batch_size):
        new_global = micro_batch_size * self._config.gradient_accumulation_steps * self.dp_world_size
        self._config.train_batch_size = new_global
        self._config.train_micro_batch_size_per_gpu = micro_batch_size

    def steps_per_print(self):
        return self._config.steps_per_print

    def zero_optimization(self):
        return self._config.zero_enabled

    def zero_optimization_stage(self):
        return self._config.zero_optimization_stage

    def zero_allow_untested_optimizer(self):
        return self._config.zero_allow_untested_optimizer

    def zero_force_ds_cpu_optimizer(self):
        return self._config.zero_force_ds_cpu_optimizer

    def zero_reduce_scatter(self):
        return self._config.zero_config.reduce_scatter

    def zero_overlap_comm(self):
        return self._config.zero_config.overlap_comm

    def zero_offload_optimizer(self):
        return self._config.zero_config.offload_optimizer

    def zero_offload_param(self):
        return self._config.zero_config.offload_param

    def zero_use_cpu_optimizer(self):
        oc = self._config.zero_config.offload_optimizer
        return oc is not None and str(getattr(oc.device, "value", oc.device)) in ("cpu", "nvme")

    def zero_cpu_offload(self):
        return self.zero_use_cpu_optimizer()

    def zero_sub_group_size(self):
        return self._config.zero_config.sub_group_size

    def zero_reduce_bucket_size(self):
        return self._config.zero_config.reduce_bucket_size

    def zero_allgather_bucket_size(self):
        return self._config.zero_config.allgather_bucket_size

    def zero_optimization_partition_gradients(self):
        return self.zero_optimization_stage() >= ZeroStageEnum.gradients

    def zero_optimization_partition_weights(self):
        return self.zero_optimization_stage() >= ZeroStageEnum.weights

    def zero_contiguous_gradients(self):
        return self._config.zero_config.contiguous_gradients

    def zero_load_from_fp32_weights(self):
        return self._config.zero_config.load_from_fp32_weights

    def zero_elastic_checkpoint(self):
        return self._config.zero_config.elastic_checkpoint

    def zero_max_live_parameters(self):
        return self._config.zero_config.max_live_parameters

    def zero_max_reuse_distance(self):
        return self._config.zero_config.max_reuse_distance

    def zero_prefetch_bucket_size(self):
        return self._config.zero_config.prefetch_bucket_size

    def zero_param_persistence_threshold(self):
        return self._config.zero_config.param_persistence_threshold

    def zero_model_persistence_threshold(self):
        return self._config.zero_config.model_persistence_threshold

    def zero_gather_16bit_weights_on_model_save(self):
        return self._config.zero_config.gather_16bit_weights_on_model_save

    def zero_grad_hooks(self):
        return True

    def zero_legacy_stage1(self):
        return self._config.zero_config.legacy_stage1

    def zero_ignore_unused_parameters(self):
        return self._config.zero_config.ignore_unused_parameters

    def zero_round_robin_gradients(self):
        return self._config.zero_config.round_robin_gradients

    def zero_hpz_partition_size(self):
        return self._config.zero_config.zero_hpz_partition_size

    def zero_quantized_weights(self):
        return self._config.zero_config.zero_quantized_weights

    def zero_quantized_nontrainable_weights(self):
        return self._config.zero_config.zero_quantized_nontrainable_weights

    def zero_quantized_gradients(self):
        return self._config.zero_config.zero_quantized_gradients

    def mics_shard_size(self):
        return self._config.mics_shard_size

    def fp16_enabled(self):
        return self._config.fp16_enabled

    def bfloat16_enabled(self):
        return self._config.bfloat16_enabled

    def fp16_master_weights_and_gradients(self):
        return self._config.fp16_master_weights_and_gradients

    def amp_enabled(self):
        return self._config.amp_enabled

    def amp_params(self):
        return self._config.amp_params

    def fp16_auto_cast(self):
        return self._config.fp16_auto_cast

    def loss_scale(self):
        return self._config.loss_scale

    def dynamic_loss_scale(self):
        return self._config.loss_scale == 0

    def initial_dynamic_scale(self):
        return self._config.initial_dynamic_scale

    def dynamic_loss_scale_args(self):
        return self._config.dynamic_loss_scale_args

    def gradient_clipping(self):
        return self._config.gradient_clipping

    def gradient_predivide_factor(self):
        return self._config.gradient_predivide_factor

    def postscale_gradients(self):
        return not self._config.prescale_gradients

    @property
    def communication_data_type(self):
        if self._config.communication_data_type is not None:
            return self._config.communication_data_type
        if self.fp16_enabled():
            return torch.float16
        if self.bfloat16_enabled():
            return torch.bfloat16
        return torch.float32

    @communication_data_type.setter
    def communication_data_type(self, value):
        self._config.communication_data_type = value

    def get_data_types(self):
        model_dtype = torch.float16 if self.fp16_enabled() else (torch.bfloat16 if self.bfloat16_enabled() else torch.float32)
        acc = self._config.grad_accum_dtype
        grad_accum_dtype = None if acc is None else {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16}[acc]
        return model_dtype, grad_accum_dtype

    def optimizer_name(self):
        return self.client_optimizer.__class__.__name__ if self.client_optimizer else self._config.optimizer_name

    def optimizer_params(self):
        return self._config.optimizer_params

    def optimizer_legacy_fusion(self):
        return self._config.optimizer_legacy_fusion

    def scheduler_name(self):
        return self._config.scheduler_name

    def scheduler_params(self):
        return self._config.scheduler_params

    def sparse_gradients_enabled(self):
        return self._config.sparse_gradients_enabled

    def wall_clock_breakdown(self):
        return self._config.wall_clock_breakdown

    def memory_breakdown(self):
        return self._config.memory_breakdown

    def flops_profiler_enabled(self):
        return self._config.flops_profiler_config.enabled or self.autotuning_enabled()

    def flops_profiler_recompute_fwd_factor(self):
        return self._config.flops_profiler_config.recompute_fwd_factor

    def flops_profiler_profile_step(self):
        step = self._config.flops_profiler_config.profile_step
        if self._config.autotuning_config.enabled:
            step = self.autotuning_start_profile_step()
        return step

    def flops_profiler_module_depth(self):
        return self._config.flops_profiler_config.module_depth

    def flops_profiler_top_modules(self):
        return self._config.flops_profiler_config.top_modules

    def flops_profiler_detailed(self):
        return False if self._config.autotuning_config.enabled else self._config.flops_profiler_config.detailed

    def flops_profiler_output_file(self):
        return self._config.flops_profiler_config.output_file

    def autotuning_enabled(self):
        return self._config.autotuning_config.enabled

    def autotuning_start_profile_step(self):
        return self._config.autotuning_config.start_profile_step

    def autotuning_end_profile_step(self):
        return self._config.autotuning_config.end_profile_step

    def autotuning_metric_path(self):
        path = self._config.autotuning_config.metric_path
        if not path:
            path = os.path.join(os.getcwd(), "autotuning_metric.json")
        return path

    def autotuning_model_info_path(self):
        path = self._config.autotuning_config.model_info_path
        if not path:
            path = os.path.join(os.getcwd(), "autotuning_model_info.json")
        return path

    def autotuning_metric(self):
        return self._config.autotuning_config.metric

    def autotuning_profile_model_info(self):
        return self.autotuning_enabled() and self._config.autotuning_config.model_info and \
            self._config.autotuning_config.model_info.get("profile", False)

    def pld_enabled(self):
        return self._config.pld_enabled

    def pld_params(self):
        return self._config.pld_params

    def pld_theta(self):
        return self.pld_params()["theta"]

    def pld_gamma(self):
        return self.pld_params()["gamma"]

    def eigenvalue_enabled(self):
        return self._config.eigenvalue_enabled

    def curriculum_enabled_legacy(self):
        return self._config.curriculum_enabled_legacy

    def curriculum_params_legacy(self):
        return self._config.curriculum_params_legacy

    def data_efficiency_enabled(self):
        return self._config.data_efficiency_enabled

    def data_efficiency_config(self):
        return self._config.data_efficiency_config

    def data_sampling_enabled(self):
        return self._config.data_efficiency_config["data_sampling"]["enabled"]

    def curriculum_learning_enabled(self):
        return self._config.data_efficiency_config["data_sampling"]["curriculum_learning"]["enabled"]

    def random_ltd_enabled(self):
        return self._config.data_efficiency_config["data_routing"]["random_ltd"]["enabled"]

    def checkpoint_tag_validation_enabled(self):
        return self._config.checkpoint_tag_validation_enabled

    def checkpoint_tag_validation_fail(self):
        return self._config.checkpoint_tag_validation_fail

    def load_universal_checkpoint(self):
        return self._config.load_universal_checkpoint

    def use_node_local_storage(self):
        return self._config.use_node_local_storage

    def elasticity_enabled(self):
        return self._config.elasticity_enabled

    def dataloader_drop_last(self):
        return self._config.dataloader_drop_last

    def aio_config(self):
        return self._config.aio_config

    def swap_tensor_config(self):
        return self._config.zero_config.offload_optimizer

    def get_batch_info(self):
        return (self.train_batch_size(), self.train_micro_batch_size_per_gpu(), self.gradient_accumulation_steps())

    def graph_harvesting(self):
        return self._config.graph_harvesting

    # ------------------------------------------------------------------ setup
    def _do_args_sanity_check(self, args):
        assert "LOCAL_RANK" in os.environ or "OMPI_COMM_WORLD_LOCAL_RANK" in os.environ or not dist.is_initialized() or True
        if hasattr(args, "deepscale_config") and args.deepscale_config is not None:
            logger.warning("************ --deepscale_config is deprecated, please use --deepspeed_config ************")
            if hasattr(args, "deepspeed_config"):
                assert args.deepspeed_config is None, "Not sure how to proceed, we were given both a deepscale_config and deepspeed_config"
            args.deepspeed_config = args.deepscale_config
            args.deepscale_config = None
        if self.config is None:
            assert hasattr(args, "deepspeed_config") and args.deepspeed_config is not None, \
                "DeepSpeed requires --deepspeed_config to specify configuration file"

    def _configure_with_arguments(self, args, mpu):
        if "OMPI_COMM_WORLD_LOCAL_RANK" in os.environ:
            os.environ["LOCAL_RANK"] = os.environ.get("OMPI_COMM_WORLD_LOCAL_RANK")
        self.local_rank = int(os.environ.get("LOCAL_RANK", 0))
        if hasattr(args, "local_rank"):
            args.local_rank = self.local_rank
        if self._config is None:
            cfg = self.config if self.config is not None else args.deepspeed_config
            self._config = DeepSpeedConfig(cfg, mpu)

    def _do_sanity_check(self):
        if self.fp16_enabled() and not get_accelerator().is_fp16_supported():
            raise ValueError("Type fp16 is not supported.")
        expected = (type(None), Optimizer, type(lambda: 0))
        assert self.client_optimizer is None or isinstance(self.client_optimizer, Optimizer) or callable(self.client_optimizer), \
            f"Client Optimizer is of unexpected type {type(self.client_optimizer)}"
        if self.client_optimizer is None and self.optimizer_name() is not None:
            name = self.optimizer_name()
            assert name in DEEPSPEED_OPTIMIZERS or hasattr(torch.optim, name), f"{name} is not a supported optimizer"

    def _set_distributed_vars(self, args):
        device_rank = args.device_rank if args is not None and hasattr(args, "device_rank") else self.local_rank
        acc = get_accelerator()
        if acc.device_name() == "cuda" and device_rank >= 0:
            acc.set_device(device_rank % max(1, acc.device_count()))
            self.device = torch.device("cuda", acc.current_device())
        else:
            self.device = torch.device("cpu")
        self.world_size = dist.get_world_size()
        self.global_rank = dist.get_rank()

    def _configure_distributed_model(self, model):
        self._set_client_model(model)
        is_zero_init_model = any(getattr(p, "ds_init_sharded", False) for p in self.module.parameters())
        if self.fp16_enabled():
            if not is_zero_init_model:
                self.module.half()
        elif self.bfloat16_enabled():
            if not is_zero_init_model:
                self.module.bfloat16()
        else:
            if any(p.dtype in (torch.half, torch.bfloat16) for p in self.module.parameters()) and not is_zero_init_model:
                pass
        if not self.dont_change_device and not is_zero_init_model:
            self.module.to(self.device)
        # MoE discovery
        from ..moe.layer import MoE
        for _, m in self.module.named_modules():
            if isinstance(m, MoE):
                self.has_moe_layers = True
                self.num_experts.append(m.num_experts)
        if self.has_moe_layers:
            for _, m in self.module.named_modules():
                if isinstance(m, MoE):
                    m.set_deepspeed_parallelism(self._config.use_data_before_expert_parallel_)
        if self.mpu is not None:
            groups.mpu = self.mpu
        self.data_parallel_group = groups._get_data_parallel_group()
        self.dp_world_size = groups._get_data_parallel_world_size()
        self.seq_data_parallel_group = groups._get_sequence_data_parallel_group()
        self.seq_dp_world_size = groups._get_sequence_data_parallel_world_size()
        self.mp_world_size = groups._get_model_parallel_world_size()
        self.expert_parallel_group = groups._get_expert_parallel_group_dict()
        self.expert_data_parallel_group = groups._get_expert_data_parallel_group_dict()
        self.sequence_parallel_size = groups._get_sequence_parallel_world_size()
        if self.sequence_parallel_size > 1:
            self.communication_data_type = self._config.seq_parallel_communication_data_type
        if not (self.amp_enabled() or is_zero_init_model) and self.zero_optimization_stage() == 0 and False:
            self._broadcast_model()

    def _set_client_model(self, model):
        modules = self.__dict__.get("_modules")
        modules["module"] = model
        self.__dict__["module"] = model

    def _broadcast_model(self):
        for p in self.module.parameters():
            if torch.is_tensor(p) and p.numel() > 0:
                dist.broadcast(p.data, groups._get_broadcast_src_rank(), group=self.seq_data_parallel_group)

    # ------------------------------------------------------------------ optimizer
    def _configure_optimizer(self, client_optimizer, model_parameters):
        if client_optimizer is None:
            if model_parameters is None:
                model_parameters = [p for p in self.module.parameters() if p.requires_grad]
            basic_optimizer = self._configure_basic_optimizer(model_parameters)
            log_dist(f"Using DeepSpeed Optimizer param name {self.optimizer_name()} as basic optimizer", ranks=[0])
        else:
            if isinstance(client_optimizer, Optimizer):
                basic_optimizer = client_optimizer
                log_dist("Using client Optimizer as basic optimizer", ranks=[0])
            else:
                basic_optimizer = client_optimizer(model_parameters if model_parameters is not None else self.module.parameters())
                log_dist("Using client callable to create basic optimizer", ranks=[0])
        self.basic_optimizer = basic_optimizer
        if self.zero_optimization() and not self.zero_allow_untested_optimizer():
            known = ("Adam", "AdamW", "FusedAdam", "DeepSpeedCPUAdam", "FusedLamb", "FusedLion", "DeepSpeedCPULion",
                     "DeepSpeedCPUAdagrad", "Adagrad", "SGD", "Lion", "OnebitAdam", "OnebitLamb", "ZeroOneAdam")
            assert type(basic_optimizer).__name__ in known, (
                "You are using an untested ZeRO Optimizer. Please add <\"zero_allow_untested_optimizer\": true> in the "
                "configuration file to use it.")
        self.optimizer = self._wrap_zero(basic_optimizer)
        log_dist(f"DeepSpeed Final Optimizer = {type(self.optimizer).__name__} over {self.optimizer_name()}", ranks=[0])

    def _configure_basic_optimizer(self, model_parameters):
        optimizer_parameters = dict(self.optimizer_params() or {})
        if "max_grad_norm" in optimizer_parameters:
            raise ValueError("'max_grad_norm' is not supported as an optimizer parameter, please switch to using the "
                             "deepspeed parameter 'gradient_clipping' see: https://www.deepspeed.ai/docs/config-json/#gradient-clipping for more details")
        name = self.optimizer_name()
        if name in [ADAM_OPTIMIZER, ADAMW_OPTIMIZER]:
            torch_adam = optimizer_parameters.pop(TORCH_ADAM_PARAM, False)
            adam_w_mode = optimizer_parameters.pop(ADAM_W_MODE, ADAM_W_MODE_DEFAULT)
            effective_adam_w_mode = name == ADAMW_OPTIMIZER or adam_w_mode
            if torch_adam:
                cls = torch.optim.AdamW if effective_adam_w_mode else torch.optim.Adam
                return cls(model_parameters, **optimizer_parameters)
            if self.zero_use_cpu_optimizer():
                from ..ops.adam import DeepSpeedCPUAdam
                return DeepSpeedCPUAdam(model_parameters, **optimizer_parameters, adamw_mode=effective_adam_w_mode)
            from ..ops.adam import FusedAdam
            return FusedAdam(model_parameters, **optimizer_parameters, adam_w_mode=effective_adam_w_mode)
        if name == ADAGRAD_OPTIMIZER:
            if self.zero_use_cpu_optimizer():
                from ..ops.adagrad import DeepSpeedCPUAdagrad
                return DeepSpeedCPUAdagrad(model_parameters, **optimizer_parameters)
            return torch.optim.Adagrad(model_parameters, **optimizer_parameters)
        if name == LAMB_OPTIMIZER:
            from ..ops.lamb import FusedLamb
            return FusedLamb(model_parameters, **optimizer_parameters)
        if name == LION_OPTIMIZER:
            if self.zero_use_cpu_optimizer():
                from ..ops.lion import DeepSpeedCPULion
                return DeepSpeedCPULion(model_parameters, **optimizer_parameters)
            from ..ops.lion import FusedLion
            return FusedLion(model_parameters, **optimizer_parameters)
        if name == ONEBIT_ADAM_OPTIMIZER:
            from .fp16.onebit.adam import OnebitAdam
            return OnebitAdam(model_parameters, self, **optimizer_parameters)
        if name == ZERO_ONE_ADAM_OPTIMIZER:
            from .fp16.onebit.zoadam import ZeroOneAdam
            return ZeroOneAdam(model_parameters, self, **optimizer_parameters)
        if name == ONEBIT_LAMB_OPTIMIZER:
            from .fp16.onebit.lamb import OnebitLamb
            return OnebitLamb(model_parameters, self, **optimizer_parameters)
        if name == SGD_OPTIMIZER:
            return torch.optim.SGD(model_parameters, **optimizer_parameters)
        torch_optimizer = getattr(torch.optim, name)
        return torch_optimizer(model_parameters, **optimizer_parameters)

    def _wrap_zero(self, basic_optimizer):
        model_dtype, grad_accum_dtype = self.get_data_types()
        stage = self.zero_optimization_stage()
        zc = self._config.zero_config
        dyn = self.fp16_enabled() and self.dynamic_loss_scale()
        common = dict(timers=self.timers, static_loss_scale=self.loss_scale() if self.fp16_enabled() else 1.0,
                      dynamic_loss_scale=dyn, dynamic_loss_args=self.dynamic_loss_scale_args(), clip_grad=self.gradient_clipping(),
                      mpu=self.mpu, communication_data_type=self.communication_data_type,
                      postscale_gradients=self.postscale_gradients(), gradient_predivide_factor=self.gradient_predivide_factor(),
                      gradient_accumulation_steps=self.gradient_accumulation_steps(), zero_config=zc, dtype=model_dtype,
                      gradient_accumulation_dtype=grad_accum_dtype,
                      grad_divisor=self.seq_dp_world_size / max(1, self.sequence_parallel_size) if self.sequence_parallel_size > 1 else None)
        if self.has_moe_layers:
            from ..moe.utils import MoEZeroOptimizer
            return MoEZeroOptimizer(self.module, basic_optimizer, stage=stage, dp_group=self.seq_data_parallel_group,
                                    expert_data_parallel_group=self.expert_data_parallel_group, **common)
        if stage == 3:
            common["aio_config"] = self.aio_config()
            if self.mics_shard_size() > 0:
                from .zero.mics import MiCS_Optimizer
                return MiCS_Optimizer(self.module, basic_optimizer, mics_shard_size=self.mics_shard_size(),
                                      dp_process_group=self.seq_data_parallel_group, **common)
            from .zero.stage3 import DeepSpeedZeroOptimizer_Stage3
            return DeepSpeedZeroOptimizer_Stage3(self.module, basic_optimizer, dp_process_group=self.seq_data_parallel_group, **common)
        from .zero.stage_1_and_2 import DeepSpeedZeroOptimizer
        return DeepSpeedZeroOptimizer(basic_optimizer, module=self.module, dp_process_group=self.seq_data_parallel_group,
                                      stage=stage, partition_grads=stage >= 2, **common)

    def _configure_lr_scheduler(self, client_lr_scheduler):
        lr_scheduler = self._scheduler_from_config(self.optimizer)
        if lr_scheduler:
            log_dist(f"DeepSpeed using configured LR scheduler = {self.scheduler_name()}", ranks=[0])
            self.lr_scheduler = lr_scheduler
        else:
            if isinstance(client_lr_scheduler, type(lambda: 0)) or (callable(client_lr_scheduler) and not hasattr(client_lr_scheduler, "step")):
                log_dist("DeepSpeed using client callable to create LR scheduler", ranks=[0])
                self.lr_scheduler = client_lr_scheduler(self.basic_optimizer)
            else:
                log_dist("DeepSpeed using client LR scheduler", ranks=[0])
                self.lr_scheduler = client_lr_scheduler
        log_dist(f"DeepSpeed LR Scheduler = {self.lr_scheduler}", ranks=[0])

    def _scheduler_from_config(self, optimizer):
        scheduler_name = self.scheduler_name()
        if scheduler_name is not None:
            if hasattr(lr_schedules, scheduler_name):
                scheduler = getattr(lr_schedules, scheduler_name)
            else:
                assert hasattr(torch.optim.lr_scheduler, scheduler_name), f"DeepSpeed does not recognize LR scheduler {scheduler_name}"
                scheduler = getattr(torch.optim.lr_scheduler, scheduler_name)
            return scheduler(self.basic_optimizer, **(self.scheduler_params() or {}))
        return None

    def _configure_checkpointing(self):
        from .checkpoint_engine.torch_checkpoint_engine import TorchCheckpointEngine
        self.checkpoint_engine = TorchCheckpointEngine()
        if self._config.nebula_config.enabled:
            from .checkpoint_engine.nebula_checkpoint_engine import NebulaCheckpointEngine
            self.checkpoint_engine = NebulaCheckpointEngine(config_params=self._config.nebula_config)
        dp_rank = groups._get_sequence_data_parallel_rank()
        rank = self.local_rank if self.use_node_local_storage() else dp_rank
        self.save_non_zero_checkpoint = (rank == 0) or self.zero_optimization_partition_weights()
        self.save_zero_checkpoint = self.optimizer is not None and hasattr(self.optimizer, "partition_count")

    # ------------------------------------------------------------------ data
    def deepspeed_io(self, dataset, batch_size=None, route="train", pin_memory=True, data_sampler=None, collate_fn=None,
                     num_local_io_workers=None):
        if not (isinstance(dataset, torch.utils.data.Dataset)):
            raise ValueError("Training data must be a torch Dataset")
        if batch_size is None:
            batch_size = self.train_micro_batch_size_per_gpu()
        if collate_fn is None:
            collate_fn = self.collate_fn
        if self.mpu is not None:
            dp_world, dp_rank = self.mpu.get_data_parallel_world_size(), self.mpu.get_data_parallel_rank()
        else:
            dp_world, dp_rank = self.dp_world_size, groups._get_data_parallel_rank()
        if data_sampler is None and route in ("predict", "eval"):
            data_sampler = torch.utils.data.DistributedSampler(dataset, num_replicas=dp_world, rank=dp_rank, shuffle=False)
        loader_config = {}
        if route == "train" and self.curriculum_learning_enabled():
            loader_config = {"curriculum_learning_enabled": True, "data_efficiency_config": self._data_efficiency_for_sampler(),
                             "data_parallel_group": self.seq_data_parallel_group, "gradient_accumulation_steps": self.gradient_accumulation_steps(),
                             "global_rank": self.global_rank}
        loader = DeepSpeedDataLoader(dataset=dataset, batch_size=batch_size, pin_memory=pin_memory, collate_fn=collate_fn,
                                     local_rank=self.local_rank, tput_timer=None, num_local_io_workers=num_local_io_workers,
                                     data_sampler=data_sampler, data_parallel_world_size=dp_world, data_parallel_rank=dp_rank,
                                     dataloader_drop_last=self.dataloader_drop_last(), deepspeed_dataloader_config=loader_config)
        if loader_config:
            self.training_data_sampler = loader.data_sampler
        return loader

    def _data_efficiency_for_sampler(self):
        """data_efficiency config in the key layout of data_pipeline/constants.py (what DeepSpeedDataSampler indexes)."""
        import copy
        cfg = copy.deepcopy(self._config.data_efficiency_config)
        cfg.setdefault("seed", 1234)
        return cfg

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        if getattr(self, "training_data_sampler", None) is not None:
            self.training_data_sampler.set_custom_curriculum_learning_schedule(schedule_func_dict)

    # ------------------------------------------------------------------ random-LTD (data routing)
    def random_ltd_config(self):
        return self._config.data_efficiency_config["data_routing"]["random_ltd"]

    def random_ltd_initialize(self):
        """Bind every RandomLayerTokenDrop wrapper (see data_routing.helper.convert_to_random_ltd) to the scheduler, in the
        order given by `random_ltd_layer_id`.  Parity: engine.py:556-577 upstream."""
        from collections import deque
        from .data_pipeline import constants as K
        from .data_pipeline.data_routing.basic_layer import RandomLayerTokenDrop
        assert self.random_ltd_enabled()
        cfg = self.random_ltd_config()
        wanted = deque(sorted(cfg[K.RANDOM_LTD_LAYER_ID]))
        count = 0
        for name, layer in self.module.named_modules():
            if isinstance(layer, RandomLayerTokenDrop):
                if len(wanted) != 0 and str(wanted[0]) in name:
                    layer.init_config(cfg, self.random_ltd_scheduler, count)
                    wanted.popleft()
                    count += 1
        if cfg[K.RANDOM_LTD_LAYER_NUM] != count:
            raise ValueError(f"random_ltd_layer_num {cfg[K.RANDOM_LTD_LAYER_NUM]} must be equivalent to the len of random_ltd_layer_id {count}")

    # ------------------------------------------------------------------ train / eval
    def train(self, mode=True):
        self.warn_unscaled_loss = True
        self.module.train(mode)

    def eval(self):
        self.warn_unscaled_loss = True
        self.module.train(False)

    def is_gradient_accumulation_boundary(self):
        if self._is_gradient_accumulation_boundary is None:
            return (self.micro_steps + 1) % self.gradient_accumulation_steps() == 0
        return self._is_gradient_accumulation_boundary

    def set_gradient_accumulation_boundary(self, is_boundary):
        self._is_gradient_accumulation_boundary = is_boundary
        if self.optimizer is not None:
            self.optimizer.is_gradient_accumulation_boundary = is_boundary

    def zero_grad(self):
        for p in self.module.parameters():
            p.grad = None

    def _scale_loss_by_gas(self, prescaled_loss):
        if isinstance(prescaled_loss, torch.Tensor):
            return prescaled_loss / self.gradient_accumulation_steps()
        if isinstance(prescaled_loss, (tuple, list)):
            return type(prescaled_loss)(l / self.gradient_accumulation_steps() if isinstance(l, torch.Tensor) else l for l in prescaled_loss)
        if self.warn_unscaled_loss:
            logger.warning(f"DeepSpeed unable to scale loss because of type: {type(prescaled_loss)}")
            self.warn_unscaled_loss = False
        return prescaled_loss

    def forward(self, *inputs, **kwargs):
        flops_profiler_active = (self.flops_profiler_enabled() and self.global_steps == self.flops_profiler_profile_step()
                                 and self.global_rank == 0)
        if flops_profiler_active:
            self.flops_profiler.start_profile(ignore_list=None)
        if self.module.training and self.random_ltd_scheduler is not None:
            self.random_ltd_scheduler.update_seq(self.global_steps)
        if self.module.training:
            if self.progressive_layer_drop:
                kwargs.update(self.progressive_layer_drop.get_state())
            if self.curriculum_enabled_legacy():
                self.curriculum_scheduler_legacy.update_difficulty(self.global_steps + 1)
                if self.curriculum_params_legacy()["curriculum_type"] == "seqlen":
                    kwargs.update({"curriculum_seqlen": self.curriculum_scheduler_legacy.get_current_difficulty()})
        if self.training_dataloader is None:
            self.tput_timer.start()
        profile_model_info = self.autotuning_profile_model_info()
        if profile_model_info:
            mem_before = get_accelerator().memory_allocated() if get_accelerator().is_available() else 0
        self._start_timers(self.engine_timers.forward_timers)
        if self.fp16_auto_cast():
            inputs = tuple(t.half() if torch.is_tensor(t) and t.is_floating_point() else t for t in inputs)
        loss = self.module(*inputs, **kwargs)
        self._stop_timers(self.engine_timers.forward_timers)
        if flops_profiler_active:
            self.flops_profiler.stop_profile()
        if profile_model_info:
            mem_after = get_accelerator().memory_allocated() if get_accelerator().is_available() else 0
            self._autotuning_write_model_info(mem_after - mem_before)
        return loss

    def allreduce_gradients(self, bucket_size=MEMORY_OPT_ALLREDUCE_SIZE, local_only=False):
        if self.optimizer is not None and hasattr(self.optimizer, "end_micro_step"):
            self.optimizer.end_micro_step(self.is_gradient_accumulation_boundary(), local_only=local_only)

    def backward(self, loss, allreduce_gradients=True, release_loss=False, retain_graph=False, scale_wrt_gas=True):
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        if self.gradient_accumulation_steps() > 1 and scale_wrt_gas:
            loss = self._scale_loss_by_gas(loss.float())
        if self.monitor.enabled and self.is_gradient_accumulation_boundary():
            self.losses = loss.detach() if self.losses is None else self.losses + loss.detach()
        self._start_timers(self.engine_timers.backward_timers)
        self._start_timers(self.engine_timers.backward_inner_timers)
        self.optimizer.backward(loss, retain_graph=retain_graph)
        self._stop_timers(self.engine_timers.backward_inner_timers)
        self._start_timers(self.engine_timers.backward_reduce_timers)
        if allreduce_gradients:
            # enable_backward_allreduce is switched off by the communication-compressed optimizers once they start
            # compressing: the micro-step bookkeeping still runs, only the dense data-parallel reduction is skipped
            self.allreduce_gradients(local_only=not self.enable_backward_allreduce)
        self._stop_timers(self.engine_timers.backward_reduce_timers)
        self._stop_timers(self.engine_timers.backward_timers)
        if release_loss:
            pass
        return loss

    # ------------------------------------------------------------------ autotuning experiment protocol
    def _autotuning_tick(self):
        """An autotuning experiment runs the user script with `autotuning.enabled`: steps [start_profile_step,
        end_profile_step] are timed (device-synchronised wall clock around whole steps), the metrics are written to
        `metric_path` and the process exits (engine.py:2204-2260 upstream)."""
        import time
        acc = get_accelerator()
        if self.global_steps == self.autotuning_start_profile_step():
            if acc.is_available():
                acc.synchronize()
            self._autotuning_t0 = time.perf_counter()
        elif self.global_steps == self.autotuning_end_profile_step() + 1 and getattr(self, "_autotuning_t0", None) is not None:
            if acc.is_available():
                acc.synchronize()
            n = self.autotuning_end_profile_step() + 1 - self.autotuning_start_profile_step()
            latency_ms = (time.perf_counter() - self._autotuning_t0) * 1e3 / max(n, 1)
            if self.global_rank == 0:
                import json
                msg = {"latency": latency_ms, "throughput": self.train_batch_size() * 1e3 / latency_ms,
                       "FLOPS_per_gpu": getattr(self, "_autotuning_flops", 0) * self.gradient_accumulation_steps() * 1e3 / latency_ms}
                path = self.autotuning_metric_path()
                os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
                with open(path, "w") as f:
                    json.dump(msg, f)
                log_dist(f"Wrote metrics to {path}, {os.path.abspath(path)}", ranks=[0])
            if dist.is_initialized():
                dist.barrier()
            import sys
            sys.exit(0)

    def _autotuning_write_model_info(self, activation_mem):
        import json
        params = list(self.module.parameters())
        info = {"num_params": sum(getattr(p, "ds_numel", p.numel()) for p in params) * self.mp_world_size,
                "trainable_num_params": sum(getattr(p, "ds_numel", p.numel()) for p in params if p.requires_grad) * self.mp_world_size,
                "hidden_size": (self._config.autotuning_config.model_info or {}).get("hidden_size"),
                "num_layers": (self._config.autotuning_config.model_info or {}).get("num_layers"),
                "activation_mem_per_gpu": activation_mem}
        if self.global_rank == 0:
            path = self.autotuning_model_info_path()
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            with open(path, "w") as f:
                json.dump(info, f)
        if dist.is_initialized():
            dist.barrier()
        import sys
        sys.exit(0)

    def clip_fp32_gradients(self):
        pass  # clipping is folded into the fused optimizer step (device-side coefficient)

    def _configure_quantization(self):
        """MoQ (quantise the 16-bit weights in the optimizer step, reference engine.py:1371-1401): enabled by the compression
        config's weight_quantization block unless it asks for quantisation in the forward pass instead."""
        (in_forward, enabled, q_groups, q_mixed_fp16, q_change_ratio, q_type, q_rounding, q_verbose, use_kernel) = self.quantize_training()
        if not enabled or in_forward:
            return None
        assert self.fp16_enabled() or self.bfloat16_enabled(), "MoQ weight quantization needs 16-bit training (fp16 or bf16)"
        from .quantize import Quantizer
        return Quantizer(q_groups, q_mixed_fp16, q_change_ratio, q_type, q_rounding, q_verbose, self.eigenvalue_enabled(), use_kernel,
                         self.eigenvalue_layer_num() if self.eigenvalue_enabled() else 0)

    def _moq_quantize(self, block_eigenvalue=None):
        """Re-quantise the marked module weights after a step.  The module's 16-bit parameters are views into the flat store,
        so they are the tensors the next forward reads; sharded (stage 3) weights are left to forward-time quantisation."""
        if self.quantizer is None or (self.compression_scheduler is not None
                                      and not getattr(self.compression_scheduler, "weight_quantization_enabled", True)):
            return
        params = [p for p in self.module.parameters() if getattr(p, "start_bits", None) and p.numel() > 0]
        overflow = bool(getattr(self.optimizer, "overflow", False)) if self.fp16_enabled() else False
        self.quantizer.quantize([params], overflow, self.eigenvalue_enabled(), block_eigenvalue)

    def _take_model_step(self, lr_kwargs, block_eigenvalue={}):
        if self.eigenvalue is not None and self.gas_boundary_ctr % max(1, self._config.eigenvalue_gas_boundary_resolution) == 0:
            # MoQ: curvature per block feeds the quantisation period of that block (engine.py:2108-2140 upstream)
            self.block_eigenvalue = self.eigenvalue.compute_eigenvalue(self.module, self.device, self.optimizer.loss_scale
                                                                       if hasattr(self.optimizer, "loss_scale") else 1.0)
            if self.compression_scheduler is not None and hasattr(self.compression_scheduler, "set_block_eigenvalue"):
                self.compression_scheduler.set_block_eigenvalue(self.block_eigenvalue)
        self.optimizer.step()
        self._moq_quantize(getattr(self, "block_eigenvalue", None) if self.eigenvalue is not None else None)
        if self.progressive_layer_drop:
            self.progressive_layer_drop.update_state(self.global_steps)
        self._global_grad_norm = self.optimizer.get_global_grad_norm()
        overflow = getattr(self.optimizer, "overflow", False)
        self._step_applied = not overflow
        if overflow:
            self.skipped_steps += 1
        else:
            if self.lr_scheduler is not None:
                try:
                    self.lr_scheduler.step(**(lr_kwargs or {}))
                except TypeError:
                    self.lr_scheduler.step()
        if hasattr(self, "compression_scheduler") and self.compression_scheduler is not None:
            self.compression_scheduler.step()
        self.global_steps += 1
        self.global_samples += self.train_batch_size()

    def step(self, lr_kwargs=None):
        assert self.optimizer is not None, "must provide optimizer during init in order to use step"
        self._start_timers(self.engine_timers.step_timers)
        report_progress = False
        self._step_applied = False
        if self.is_gradient_accumulation_boundary():
            self.gas_boundary_ctr += 1
            self._take_model_step(lr_kwargs)
            report_progress = self.global_rank == 0 if self.global_rank else True
        self.tput_timer.stop(global_step=self.is_gradient_accumulation_boundary(), report_speed=report_progress)
        self._stop_timers(self.engine_timers.step_timers)
        if self.flops_profiler is not None and self.flops_profiler.started and self.is_gradient_accumulation_boundary():
            if self.autotuning_enabled():
                self._autotuning_flops = self.flops_profiler.get_total_flops() * 3  # fwd + 2x for bwd
            elif self.global_rank == 0:
                self.flops_profiler.print_model_profile(profile_step=self.global_steps - 1, module_depth=self.flops_profiler_module_depth(),
                                                        top_modules=self.flops_profiler_top_modules(), detailed=self.flops_profiler_detailed(),
                                                        output_file=self.flops_profiler_output_file())
            self.flops_profiler.end_profile()
        if self.autotuning_enabled() and self.is_gradient_accumulation_boundary():
            self._autotuning_tick()
        if self.is_gradient_accumulation_boundary():
            if self.monitor.enabled and self.global_rank == 0:
                events = [("Train/Samples/lr", self.get_lr()[0], self.global_samples)]
                if self.losses is not None:
                    events.append(("Train/Samples/train_loss", float(self.losses), self.global_samples))
                    self.losses = None
                if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                    events.append(("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples))
                self.monitor.write_events(events)
            if self.steps_per_print() and self.global_steps % self.steps_per_print() == 0 and self.global_rank == 0:
                self._report_progress(self.global_steps)
            if self.wall_clock_breakdown() and self.steps_per_print() and self.global_steps % self.steps_per_print() == 0:
                self.timers.log(self.engine_timers.global_timers, normalizer=self.steps_per_print())
        if self.wall_clock_breakdown():
            pass
        self.micro_steps += 1

    def _start_timers(self, names):
        for n in names:
            self.timers(n).start()

    def _stop_timers(self, names):
        record = self.is_gradient_accumulation_boundary()
        for n in names:
            self.timers(n).stop(record=record)

    def _report_progress(self, step):
        lr = self.get_lr()
        mom = self.get_mom()
        log_dist(f"step={step}, skipped={self.skipped_steps}, lr={lr}, mom={mom}", ranks=[0])

    def was_step_applied(self):
        return self._step_applied

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def get_type(self):
        return [g.get("type", None) for g in self.optimizer.param_groups]

    def get_mom(self):
        out = []
        for g in self.optimizer.param_groups:
            out.append(g["betas"] if "betas" in g else g.get("momentum", None))
        return out

    def get_pld_theta(self):
        return self.progressive_layer_drop.get_theta() if self.progressive_layer_drop else None

    def get_global_grad_norm(self):
        n = self._global_grad_norm
        return float(n.item()) if torch.is_tensor(n) else n

    @property
    def cur_scale(self):
        return self.optimizer.cur_scale if self.optimizer is not None else 1.0

    def empty_partition_cache(self):
        if hasattr(self.optimizer, "empty_partition_cache"):
            self.optimizer.empty_partition_cache()
            get_accelerator().empty_cache()

    def destroy(self):
        if self.optimizer is not None and hasattr(self.optimizer, "_hook_handles"):
            for h in self.optimizer._hook_handles:
                h.remove()
            for h in getattr(self.optimizer, "_module_hooks", []):
                h.remove()

    # ------------------------------------------------------------------ state dicts
    def module_state_dict(self, destination=None, prefix="", keep_vars=False, exclude_frozen_parameters=False):
        sd = self.module.state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars)
        if exclude_frozen_parameters:
            for n, p in self.module.named_parameters():
                if not p.requires_grad and n in sd:
                    del sd[n]
        return sd

    def load_module_state_dict(self, checkpoint=None, strict=True, custom_load_fn=None, fetch_z3_params=False, state_dict=None):
        module_state_dict = state_dict if state_dict is not None else checkpoint["module"]
        if custom_load_fn:
            custom_load_fn(src=module_state_dict, dst=self.module)
            return
        z3 = self.zero_optimization_partition_weights() and self.optimizer is not None and self.optimizer.shard_world > 1
        if z3:
            for u in self.optimizer.units:
                with self.optimizer.gathered_unit(u):
                    for p in u.params:
                        name = self.param_names.get(p)
                        src = module_state_dict.get(name) if name is not None else None
                        if src is None or src.numel() == 0:
                            continue  # ZeRO-3 model-states files hold empty placeholders for partitioned parameters
                        if src.numel() != p.data.numel():
                            raise RuntimeError(f"checkpoint tensor '{name}' has {src.numel()} elements, the parameter has {p.data.numel()}")
                        p.data.copy_(src.to(p.device, p.dtype).view_as(p.data))
                    self.optimizer._write_back_unit(u)
            # buffers
            for n, b in self.module.named_buffers():
                if n in module_state_dict:
                    b.data.copy_(module_state_dict[n])
        else:
            self.module.load_state_dict(module_state_dict, strict=strict)
            if self.optimizer is not None and hasattr(self.optimizer, "refresh_from_params"):
                self.optimizer.refresh_from_params()

    def _get_zero_param_shapes(self):
        shapes = OrderedDict()
        for u in self.optimizer.units:
            for p in u.params:
                shapes[self.param_names.get(p, str(p.ds_id))] = torch.Size(p.ds_shape)
        return [shapes]

    def _zero3_consolidated_16bit_state_dict(self):
        """Gather every unit (one at a time) and return the full 16-bit state dict on rank 0."""
        if not self.zero_optimization_partition_weights():
            raise ValueError("this function requires ZeRO-3 mode")
        state_dict = OrderedDict() if dist.get_rank() == 0 else None
        for u in self.optimizer.units:
            with self.optimizer.gathered_unit(u):
                if dist.get_rank() == 0:
                    for p in u.params:
                        state_dict[self.param_names.get(p)] = p.detach().cpu().clone()
        if dist.get_rank() == 0:
            for n, b in self.module.named_buffers():
                state_dict[n] = b.detach().cpu()
        return state_dict

    def save_fp16_model(self, save_dir, save_filename="pytorch_model.bin"):
        return self.save_16bit_model(save_dir, save_filename)

    def save_16bit_model(self, save_dir, save_filename="pytorch_model.bin", exclude_frozen_parameters=False):
        path = os.path.join(save_dir, save_filename)
        if self.zero_optimization_partition_weights():
            if self.zero_gather_16bit_weights_on_model_save():
                state_dict = self._zero3_consolidated_16bit_state_dict()
            else:
                logger.info(f"Did not save the model {path} because `stage3_gather_16bit_weights_on_model_save` is False")
                return False
        else:
            state_dict = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
        tag = f"global_step{self.global_steps}"
        tag = str(tag)
        self.checkpoint_engine.create(tag)
        if dist.get_rank() == 0:
            self.checkpoint_engine.makedirs(save_dir, exist_ok=True)
            logger.info(f"Saving model weights to {path}, tag: {tag}")
            self.checkpoint_engine.save(state_dict, path)
        self.checkpoint_engine.commit(tag)
        return True

    # ------------------------------------------------------------------ checkpointing (file layout of the reference)
    def _get_ckpt_name(self, checkpoints_path, tag, mp_placeholder=None):
        mp_rank_str = mp_placeholder if mp_placeholder is not None else f"{0 if self.mpu is None else self.mpu.get_model_parallel_rank():02d}"
        if self.zero_optimization_partition_weights():
            filename = "zero_pp_rank_{}".format(groups._get_sequence_data_parallel_rank())
            return os.path.join(checkpoints_path, str(tag), f"{filename}_mp_rank_{mp_rank_str}_model_states.pt")
        return os.path.join(checkpoints_path, str(tag), "mp_rank_" + mp_rank_str + "_model_states.pt")

    def _get_zero_ckpt_prefix(self, dp_rank, bf16_mode):
        return f'{"bf16_" if bf16_mode else ""}zero_pp_rank_{dp_rank}'

    def _get_rank_zero_ckpt_name(self, checkpoints_path, tag, mp_rank, dp_rank, bf16_mode):
        file_prefix = self._get_zero_ckpt_prefix(dp_rank, bf16_mode=bf16_mode)
        return os.path.join(checkpoints_path, str(tag), f"{file_prefix}_mp_rank_{mp_rank:02d}_optim_states.pt")

    def _get_zero_ckpt_name(self, checkpoints_path, tag):
        mp_rank = 0 if self.mpu is None else self.mpu.get_model_parallel_rank()
        pp_rank = groups._get_sequence_data_parallel_rank()
        bf16_mode = self.bfloat16_enabled()
        return self._get_rank_zero_ckpt_name(checkpoints_path, tag, mp_rank, pp_rank, bf16_mode)

    def _get_expert_ckpt_name(self, checkpoints_path, layer_id, expert_id, tag, mp_rank=0):
        return os.path.join(checkpoints_path, str(tag), f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")

    def _get_optimizer_ckpt_name(self, checkpoints_path, tag, expp_rank):
        mp_rank = 0 if self.mpu is None else self.mpu.get_model_parallel_rank()
        return os.path.join(checkpoints_path, str(tag), f"expp_rank_{expp_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")

    def _checkpoint_tag_validation(self, tag):
        if self.checkpoint_tag_validation_enabled() and dist.is_initialized() and dist.get_world_size() > 1:
            s_hash = hashlib.sha1(tag.encode())
            bhash = torch.ByteTensor([s_hash.digest()]).flatten().to(self.device)
            max_bhash, min_bhash = bhash.clone(), bhash.clone()
            dist.all_reduce(max_bhash, op=dist.ReduceOp.MAX)
            dist.all_reduce(min_bhash, op=dist.ReduceOp.MIN)
            valid = all(min_bhash == bhash) and all(max_bhash == bhash)
            msg = (f"[rank={dist.get_rank()}] The checkpoint tag name '{tag}' is not consistent across all ranks. "
                   "Including rank unique information in checkpoint tag could cause issues when restoring with different world sizes.")
            if self.checkpoint_tag_validation_fail():
                assert valid, msg
            elif not valid:
                logger.warning(msg)

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True, exclude_frozen_parameters=False):
        if self.optimizer is not None and hasattr(self.optimizer, "checkpoint_event_prologue"):
            self.optimizer.checkpoint_event_prologue()
        rank = self.local_rank if self.use_node_local_storage() else self.global_rank
        if rank == 0:
            self.checkpoint_engine.makedirs(save_dir, exist_ok=True)
        dist.barrier()
        if tag is None:
            tag = f"global_step{self.global_steps}"
        tag = str(tag)
        self.checkpoint_engine.create(tag)
        self._checkpoint_tag_validation(tag)
        if self.has_moe_layers:
            self.save_non_zero_checkpoint = False
            self._create_checkpoint_file(save_dir, tag, False)
            self._save_moe_checkpoint(save_dir, tag, client_state=client_state, exclude_frozen_parameters=exclude_frozen_parameters)
        if self.save_non_zero_checkpoint:
            self._create_checkpoint_file(save_dir, tag, False)
            self._save_checkpoint(save_dir, tag, client_state=client_state, exclude_frozen_parameters=exclude_frozen_parameters)
        if self.save_zero_checkpoint:
            self._create_zero_checkpoint_files(save_dir, tag)
            self._save_zero_checkpoint(save_dir, tag)
        if self.optimizer is not None and hasattr(self.optimizer, "checkpoint_event_epilogue"):
            self.optimizer.checkpoint_event_epilogue()
        self.checkpoint_engine.commit(tag)
        if save_latest and rank == 0:
            with open(os.path.join(save_dir, "latest"), "w") as fd:
                fd.write(tag)
        dist.barrier()
        return True

    def _create_checkpoint_file(self, save_dir, tag, zero_checkpoint):
        name_function = self._get_zero_ckpt_name if zero_checkpoint else self._get_ckpt_name
        try:
            checkpoint_name = name_function(save_dir, tag)
            self.checkpoint_engine.makedirs(os.path.dirname(checkpoint_name), exist_ok=True)
        except Exception:
            logger.error(f"Failed saving model checkpoint to {save_dir} with tag {tag}")
            return False
        return True

    def _create_zero_checkpoint_files(self, save_dir, tag):
        success = True
        for rank in range(dist.get_world_size(self.optimizer.dp_group) if dist.is_initialized() else 1):
            if rank == self.global_rank:
                success = self._create_checkpoint_file(save_dir, tag, True)
            dist.barrier(group=self.optimizer.dp_group)
        return success

    def _get_shared_params(self):
        shared_index, shared = {}, OrderedDict()
        def get_layer_state_dict(module, prefix=""):
            for name, param in module.named_parameters(recurse=False):
                if param is None or not hasattr(param, "ds_id"):
                    continue
                key = prefix + name
                if param.ds_id in shared_index:
                    shared[key] = shared_index[param.ds_id]
                else:
                    shared_index[param.ds_id] = key
            for name, child in module.named_children():
                if child is not None:
                    get_layer_state_dict(child, prefix + name + ".")
        get_layer_state_dict(self.module, prefix="")
        return shared

    def _save_checkpoint(self, save_dir, tag, client_state={}, exclude_frozen_parameters=False):
        save_path = self._get_ckpt_name(save_dir, tag)
        zero_optimizer_state = self.optimizer is not None and hasattr(self.optimizer, "partition_count")
        z3 = self.zero_optimization_partition_weights() and self.optimizer is not None and self.optimizer.shard_world > 1
        if z3:
            module_sd = OrderedDict((n, torch.empty(0)) for n, _ in self.module.named_parameters())
            for n, b in self.module.named_buffers():
                module_sd[n] = b.detach().cpu()
        else:
            module_sd = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
        state = dict(module=module_sd,
                     buffer_names=[n for n, _ in self.module.named_buffers()],
                     optimizer=None if zero_optimizer_state else (self.optimizer.state_dict() if self.optimizer else None),
                     param_shapes=self._get_zero_param_shapes() if zero_optimizer_state else None,
                     frozen_param_shapes=None, frozen_param_fragments=None,
                     shared_params=self._get_shared_params() if zero_optimizer_state else None,
                     lr_scheduler=self.lr_scheduler.state_dict() if self.lr_scheduler is not None else None,
                     data_sampler=(self.training_data_sampler.state_dict() if getattr(self, "training_data_sampler", None) is not None and
                                   hasattr(self.training_data_sampler, "state_dict") else None),
                     random_ltd=self.random_ltd_scheduler.state_dict() if self.random_ltd_scheduler is not None else None,
                     sparse_tensor_module_names=[],
                     skipped_steps=self.skipped_steps, global_steps=self.global_steps, global_samples=self.global_samples,
                     dp_world_size=self.seq_dp_world_size, mp_world_size=self.mp_world_size, ds_config=self._config._param_dict,
                     ds_version=__version__)
        if z3:
            state["ds_lp_shard"] = self.optimizer.lp_shard.detach().cpu()
        state.update(client_state)
        log_dist(message=f"Saving model checkpoint: {save_path}", ranks=[0])
        self.checkpoint_engine.save(state, save_path)

    def _save_zero_checkpoint(self, save_path, tag):
        zero_checkpoint_name = self._get_zero_ckpt_name(save_path, tag)
        zero_sd = dict(optimizer_state_dict=self.optimizer.state_dict(), ds_config=self._config._param_dict, ds_version=__version__)
        self.checkpoint_engine.save(zero_sd, zero_checkpoint_name)
        if self.global_rank == 0:
            self._copy_recovery_script(save_path)
        logger.info(f"zero checkpoint saved {zero_checkpoint_name}")

    def _copy_recovery_script(self, save_path):
        import shutil
        import stat
        base_dir = os.path.dirname(os.path.dirname(__file__))
        script = "zero_to_fp32.py"
        src = os.path.join(base_dir, "utils", script)
        dst = os.path.join(save_path, script)
        shutil.copyfile(src, dst)
        os.chmod(dst, os.stat(dst).st_mode | stat.S_IEXEC)

    def _save_moe_checkpoint(self, save_dir, tag, client_state={}, exclude_frozen_parameters=False):
        from ..moe.utils import save_moe_checkpoint
        save_moe_checkpoint(self, save_dir, tag, client_state, exclude_frozen_parameters)

    def load_checkpoint(self, load_dir, tag=None, load_module_strict=True, load_optimizer_states=True, load_lr_scheduler_states=True,
                        load_module_only=False, custom_load_fn=None):
        if tag is None:
            latest_tag = "latest_universal" if self.load_universal_checkpoint() else "latest"
            latest_path = os.path.join(load_dir, latest_tag)
            if os.path.isfile(latest_path):
                with open(latest_path, "r") as fd:
                    tag = fd.read().strip()
            else:
                if self.load_universal_checkpoint():
                    raise ValueError(f"Invalid for universal checkpoint: {latest_path} does not exist")
                logger.warning(f"Unable to find latest file at {latest_path}, if trying to load latest checkpoint please "
                               "ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
                return None, None
        if self.load_universal_checkpoint():
            from ..checkpoint.universal_checkpoint import load_universal_into_engine
            return load_universal_into_engine(self, load_dir, tag)
        load_path, client_states = self._load_checkpoint(load_dir, tag, load_module_strict=load_module_strict,
                                                         load_optimizer_states=load_optimizer_states,
                                                         load_lr_scheduler_states=load_lr_scheduler_states,
                                                         load_module_only=load_module_only, custom_load_fn=custom_load_fn)
        load_zero_checkpoint = load_path is not None and self.optimizer is not None and hasattr(self.optimizer, "partition_count") and not load_module_only
        if load_zero_checkpoint:
            if load_optimizer_states:
                success = self._load_zero_checkpoint(load_dir, tag, load_optimizer_states=load_optimizer_states)
                if not success:
                    self.optimizer.refresh_from_params() if hasattr(self.optimizer, "refresh_from_params") else None
        return load_path, client_states

    def _load_checkpoint(self, load_dir, tag, load_module_strict=True, load_optimizer_states=True, load_lr_scheduler_states=True,
                         load_module_only=False, custom_load_fn=None):
        ckpt_name = self._get_ckpt_name(load_dir, tag)
        if not os.path.exists(ckpt_name):
            # a checkpoint written under a different ZeRO stage uses the other file naming
            alt = os.path.join(load_dir, str(tag), "mp_rank_00_model_states.pt")
            if os.path.exists(alt):
                ckpt_name = alt
            else:
                logger.warning(f"Client provided checkpoint load path: {ckpt_name} does not exist")
                return None, None
        checkpoint = self.checkpoint_engine.load(ckpt_name, map_location="cpu")
        if self.has_moe_layers:
            from ..moe.utils import load_moe_state_dict
            load_moe_state_dict(self, load_dir, tag, checkpoint)
        z3 = self.zero_optimization_partition_weights() and self.optimizer is not None and self.optimizer.shard_world > 1
        if z3 and "ds_lp_shard" in checkpoint:
            saved_world = checkpoint.get("dp_world_size")
            if checkpoint["ds_lp_shard"].numel() != self.optimizer.lp_shard.numel():
                raise RuntimeError(
                    f"ZeRO-3 checkpoint was written with data-parallel world size {saved_world} (shard of {checkpoint['ds_lp_shard'].numel()} "
                    f"elements); this job has world size {self.optimizer.shard_world} (shard of {self.optimizer.lp_shard.numel()}).  Convert it with "
                    "checkpoint/ds_to_universal.py (loads at any world size) or utils/zero_to_fp32.py.")
            self.optimizer.lp_shard.copy_(checkpoint["ds_lp_shard"].to(self.optimizer.lp_shard.device))
            self.optimizer._refresh_master(0, self.optimizer.total_shard)
            self.optimizer._post_load_sync()
            for n, b in self.module.named_buffers():
                if n in checkpoint["module"]:
                    b.data.copy_(checkpoint["module"][n])
        else:
            self.load_module_state_dict(checkpoint=checkpoint, strict=load_module_strict, custom_load_fn=custom_load_fn)
        self.loaded_checkpoint_dp_world_size = checkpoint.get("dp_world_size")
        self.loaded_checkpoint_mp_world_size = checkpoint.get("mp_world_size")
        if load_module_only:
            deepspeed_states = ["module"]
        else:
            if self.optimizer is not None and not hasattr(self.optimizer, "partition_count") and load_optimizer_states and checkpoint.get("optimizer"):
                self.optimizer.load_state_dict(checkpoint["optimizer"])
            if load_lr_scheduler_states and self.lr_scheduler is not None and checkpoint.get("lr_scheduler") is not None:
                self.lr_scheduler.load_state_dict(checkpoint["lr_scheduler"])
            self.global_steps = checkpoint.get("global_steps", 0)
            self.global_samples = checkpoint.get("global_samples", self.global_steps * self.train_batch_size())
            self.skipped_steps = checkpoint.get("skipped_steps", 0)
            if self.random_ltd_scheduler is not None and checkpoint.get("random_ltd") is not None:
                self.random_ltd_scheduler.load_state_dict(checkpoint["random_ltd"])
            if getattr(self, "training_data_sampler", None) is not None and checkpoint.get("data_sampler") is not None:
                self.training_data_sampler.load_state_dict(checkpoint["data_sampler"])
            deepspeed_states = ["module", "sparse_tensor_module_names", "skipped_steps", "global_steps", "dp_world_size",
                                "mp_world_size", "data_sampler", "random_ltd"]
        client_state = {}
        if load_lr_scheduler_states:
            deepspeed_states.append("lr_scheduler")
        if load_optimizer_states:
            deepspeed_states.append("optimizer")
        deepspeed_states += ["buffer_names", "param_shapes", "frozen_param_shapes", "frozen_param_fragments", "shared_params",
                             "global_samples", "ds_config", "ds_version", "ds_lp_shard"]
        client_state = {key: value for key, value in checkpoint.items() if key not in deepspeed_states}
        return load_dir, client_state

    def _load_zero_checkpoint(self, load_dir, tag, load_optimizer_states=True):
        name = self._get_zero_ckpt_name(load_dir, tag)
        if not os.path.exists(name):
            logger.warning(f"The following zero checkpoint path is missing: {name}")
            return False
        if self.loaded_checkpoint_dp_world_size is not None and self.loaded_checkpoint_dp_world_size != self.seq_dp_world_size:
            raise ValueError(f"The checkpoint being loaded used a DP world size of {self.loaded_checkpoint_dp_world_size} but the "
                             f"current world size is {self.seq_dp_world_size}. Automatic adjustment of ZeRO's optimizer state "
                             "partitioning with a new world size is not currently supported; convert with ds_to_universal.")
        sd = self.checkpoint_engine.load(name, map_location="cpu")
        self.optimizer.load_state_dict(sd["optimizer_state_dict"], load_optimizer_states=load_optimizer_states,
                                       load_from_fp32_weights=self.zero_load_from_fp32_weights())
        logger.info(f"loaded zero checkpoint {name}")
        return True
