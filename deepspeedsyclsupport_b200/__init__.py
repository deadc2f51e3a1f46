"""deepspeedsyclsupport_b200 -- a Blackwell-native (B200, sm_100a) ZeRO training and inference engine with the public
API of DeepSpeed 0.12.7 (`delock/DeepSpeedSYCLSupport`): `initialize()`, `init_inference()`, `init_distributed()`,
`add_config_arguments()`, `zero.Init`, `comm.*`, `accelerator.get_accelerator()`, the ds_config schema and the
checkpoint file layout.  Usage mirrors the reference:

    import deepspeedsyclsupport_b200 as deepspeed
    engine, optimizer, _, scheduler = deepspeed.initialize(model=model, config=ds_config)

Parity: deepspeed/__init__.py:64-344.
"""
import sys
import types
from typing import Callable, Optional, Union

import torch
from torch.nn import Module
from torch.optim import Optimizer
from torch.optim.lr_scheduler import _LRScheduler

from . import comm, ops, utils  # noqa: F401
from .accelerator import get_accelerator, set_accelerator  # noqa: F401
from .comm.comm import init_distributed  # noqa: F401
from .runtime import zero  # noqa: F401
from .runtime.config import DeepSpeedConfig, DeepSpeedConfigError  # noqa: F401
from .runtime.engine import ADAM_OPTIMIZER, LAMB_OPTIMIZER, DeepSpeedEngine, DeepSpeedOptimizerCallable, DeepSpeedSchedulerCallable  # noqa: F401
from .utils import groups  # noqa: F401
from .utils.logging import log_dist, logger  # noqa: F401
from .version import __version__, __version_major__, __version_minor__, __version_patch__  # noqa: F401

__git_hash__ = None
__git_branch__ = None
dist = comm


def _lazy(name):
    import importlib
    return importlib.import_module(f"{__name__}.{name}")


def __getattr__(name):
    # heavier sub-packages load on first use: deepspeed.pipe, .moe, .sequence, .checkpointing, .inference, .module_inject ...
    lazy = {"pipe": "pipe", "moe": "moe", "sequence": "sequence", "inference": "inference", "module_inject": "module_inject",
            "profiling": "profiling", "compression": "compression", "elasticity": "elasticity", "launcher": "launcher",
            "monitor": "monitor", "autotuning": "autotuning", "checkpoint": "checkpoint", "models": "models",
            "parallel": "parallel", "model_implementations": "model_implementations", "env_report": "env_report"}
    if name in lazy:
        return _lazy(lazy[name])
    if name == "checkpointing":
        return _lazy("runtime.activation_checkpointing.checkpointing")
    if name == "PipelineModule":
        return _lazy("pipe").PipelineModule
    if name == "PipelineEngine":
        return _lazy("runtime.pipe.engine").PipelineEngine
    if name == "InferenceEngine":
        return _lazy("inference.engine").InferenceEngine
    if name == "DeepSpeedInferenceConfig":
        return _lazy("inference.config").DeepSpeedInferenceConfig
    if name == "DeepSpeedHybridEngine":
        return _lazy("runtime.hybrid_engine").DeepSpeedHybridEngine
    if name in ("DeepSpeedTransformerLayer", "DeepSpeedTransformerConfig"):
        return getattr(_lazy("ops.transformer"), name)
    if name == "OnDevice":
        return _lazy("utils.init_on_device").OnDevice
    if name == "add_tuning_arguments":
        return _lazy("runtime.lr_schedules").add_tuning_arguments
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def initialize(args=None, model: torch.nn.Module = None, optimizer: Optional[Union[Optimizer, Callable]] = None,
               model_parameters=None, training_data=None, lr_scheduler: Optional[Union[_LRScheduler, Callable]] = None,
               mpu=None, dist_init_required: Optional[bool] = None, collate_fn=None, config=None, config_params=None):
    """Build the training engine.  Returns (engine, optimizer, training_dataloader, lr_scheduler) like the reference
    (deepspeed/__init__.py:64-202); picks PipelineEngine for a PipelineModule and DeepSpeedHybridEngine when the
    `hybrid_engine` block is enabled."""
    log_dist("DeepSpeed-B200 info: version={}".format(__version__), ranks=[0])
    zero.partition_parameters.shutdown_init_context()
    assert model is not None, "deepspeed.initialize requires a model"
    try:
        init_distributed(dist_backend=get_accelerator().communication_backend_name(), dist_init_required=dist_init_required)
        if config is None:
            config = config_params
        if hasattr(args, "deepscale_config") and args.deepscale_config is not None:
            args.deepspeed_config = args.deepscale_config
            args.deepscale_config = None
        if hasattr(args, "deepspeed_config") and args.deepspeed_config is not None:
            assert config is None, "Not sure how to proceed, we were given deepspeed configs in the deepspeed arguments and deepspeed.initialize() function call"
            config = args.deepspeed_config
        assert config is not None, "DeepSpeed requires --deepspeed_config to specify configuration file"
        from .runtime.pipe.module import PipelineModule
        if not isinstance(model, PipelineModule):
            config_class = DeepSpeedConfig(config, mpu)
            if config_class.hybrid_engine.enabled:
                from .runtime.hybrid_engine import DeepSpeedHybridEngine
                engine = DeepSpeedHybridEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                               training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                               dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                               config_class=config_class)
            else:
                engine = DeepSpeedEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                         training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                         dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                         config_class=config_class)
        else:
            assert mpu is None, "mpu must be None with pipeline parallelism"
            mpu = model.mpu()
            config_class = DeepSpeedConfig(config, mpu)
            from .runtime.pipe.engine import PipelineEngine
            engine = PipelineEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                    training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                    dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                    config_class=config_class)
    finally:
        zero.partition_parameters.restore_init_context()
    return engine, engine.optimizer, engine.training_dataloader, engine.lr_scheduler


def _add_core_arguments(parser):
    group = parser.add_argument_group("DeepSpeed", "DeepSpeed configurations")
    group.add_argument("--deepspeed", default=False, action="store_true",
                       help="Enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepspeed_config", default=None, type=str, help="DeepSpeed json configuration file.")
    group.add_argument("--deepscale", default=False, action="store_true",
                       help="Deprecated enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepscale_config", default=None, type=str, help="Deprecated DeepSpeed json configuration file.")
    group.add_argument("--deepspeed_mpi", default=False, action="store_true",
                       help="Run via MPI, this will attempt to discover the necessary variables to initialize torch distributed from the MPI environment")
    return parser


def add_config_arguments(parser):
    """Add the DeepSpeed command-line arguments (--deepspeed, --deepspeed_config, ...) to an argparse parser."""
    return _add_core_arguments(parser)


def default_inference_config():
    from .inference.config import DeepSpeedInferenceConfig
    return DeepSpeedInferenceConfig().dict()


def init_inference(model, config=None, **kwargs):
    """Build an InferenceEngine (AutoTP / kernel injection / CUDA graphs).  Parity: deepspeed/__init__.py:269-344."""
    log_dist("DeepSpeed-B200 info: version={}".format(__version__), ranks=[0])
    from .inference.config import DeepSpeedInferenceConfig
    from .inference.engine import InferenceEngine
    if config is None:
        config = {}
    if isinstance(config, str):
        import json
        with open(config, "r") as f:
            config_dict = json.load(f)
    elif isinstance(config, dict):
        config_dict = config
    else:
        raise ValueError(f"'config' argument expected string or dictionary, got {type(config)}")
    overlap_keys = set(config_dict.keys()).intersection(kwargs.keys())
    for key in overlap_keys:
        if config_dict[key] != kwargs[key]:
            raise ValueError(f"Conflicting argument '{key}' in 'config':{config_dict[key]} and kwargs:{kwargs[key]}")
    config_dict.update(kwargs)
    ds_inference_config = DeepSpeedInferenceConfig(**config_dict)
    return InferenceEngine(model, config=ds_inference_config)
