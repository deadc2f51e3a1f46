"""`DeepSpeedZeRoOffload`: ZeRO stage-3 parameter sharding *without* an optimizer -- ZeRO-Inference, evaluation of a model that
does not fit on one GPU, or the forward-only half of a larger system (parity: deepspeed/runtime/zero/parameter_offload.py:201-540).

In this design one class, the flat unit store (`sharded_optimizer.ZeroShardedOptimizer`), owns sharding, fetch / release hooks,
trace-ordered prefetch and the host / NVMe parameter tiers; with `init_optimizer=None` it creates no optimizer state.  This module
exposes that configuration under the reference's name and surface (hook functions, coordinator access, persistent-parameter
marking), plus the helpers other code imports from here."""
from collections import OrderedDict

import torch

from .config import DeepSpeedZeroConfig
from .stage3 import DeepSpeedZeroOptimizer_Stage3

FWD_MODULE_STACK = list()


def is_builtin_type(obj):
    return obj.__class__.__module__ in ("__builtin__", "builtins")


def isinstance_namedtuple(obj: object) -> bool:
    """namedtuples have to be rebuilt field by field (`type(obj)(*values)`), unlike plain tuples"""
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


def _apply_to_tensors_only(module, functional, backward_function, outputs):
    """Apply `functional.apply(module, backward_function, t)` to every tensor inside a nested output structure."""
    if isinstance(outputs, (tuple, list)):
        touched = [_apply_to_tensors_only(module, functional, backward_function, o) for o in outputs]
        if isinstance_namedtuple(outputs):
            return outputs.__class__(*touched)
        return outputs.__class__(touched)
    if isinstance(outputs, dict):
        return outputs.__class__((k, _apply_to_tensors_only(module, functional, backward_function, v)) for k, v in outputs.items())
    if isinstance(outputs, torch.Tensor):
        touched = functional.apply(module, backward_function, outputs)
        if not outputs.is_leaf and isinstance(outputs, torch.nn.Parameter):
            touched.ds_param_alias = outputs
        return touched
    return outputs


class ZeROOrderedDict(OrderedDict):
    """Parameter dict of a module that gathers a sharded parameter on first access (for code that reads `module._parameters`
    directly, bypassing the module hooks)."""

    def __init__(self, parent_module, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._parent_module = parent_module
        self._in_forward = False

    def __getitem__(self, key):
        param = super().__getitem__(key)
        if param is None or not hasattr(param, "ds_unit"):
            return param
        if param.ds_unit.status.name == "NOT_AVAILABLE" and self._in_forward:
            param.all_gather()
        return param


class DeepSpeedZeRoOffload(object):

    def __init__(self, module, timers=None, ds_config=None, overlap_comm=True, prefetch_bucket_size=50000000, max_reuse_distance=1000000000,
                 max_live_parameters=1000000000, param_persistence_threshold=100000, model_persistence_threshold=None,
                 offload_param_config=None, mpu=None, zero_param_parallel_group=None, zero_quantized_weights=False,
                 zero_quantized_nontrainable_weights=False, dtype=None, dp_process_group=None, aio_config=None):
        self.module = module
        self.timers = timers
        self.dtype = dtype or next(module.parameters()).dtype
        zc = getattr(ds_config, "zero_config", None)
        if zc is None:
            zc = DeepSpeedZeroConfig(stage=3, overlap_comm=overlap_comm, offload_param=offload_param_config,
                                     stage3_prefetch_bucket_size=prefetch_bucket_size, stage3_max_live_parameters=max_live_parameters,
                                     stage3_max_reuse_distance=max_reuse_distance, stage3_param_persistence_threshold=param_persistence_threshold,
                                     zero_quantized_weights=zero_quantized_weights)
        self.offload_device = zc.offload_param.device if zc.offload_param is not None else None
        self.offload_param_pin_memory = bool(zc.offload_param.pin_memory) if zc.offload_param is not None else False
        self._store = DeepSpeedZeroOptimizer_Stage3(module, None, timers=timers, zero_config=zc, dtype=self.dtype, mpu=mpu,
                                                    dp_process_group=dp_process_group, shard_group=zero_param_parallel_group,
                                                    aio_config=aio_config)
        self.persistent_parameters = [p for u in self._store.units if u.persistent for p in u.params]
        self.forward_hooks = self.backward_hooks = self._store._module_hooks

    # ---- reference surface
    @property
    def param_store(self):
        return self._store

    def partition_all_parameters(self):
        """Release every gathered, non-persistent unit (each rank keeps only its shard)."""
        self._store.empty_partition_cache()

    def empty_partition_cache(self):
        self.partition_all_parameters()

    def get_param_coordinator(self, training):
        from .partitioned_param_coordinator import PartitionedParameterCoordinator
        return PartitionedParameterCoordinator.for_store(self._store, training)

    def mark_persistent_parameters(self, param_threshold, model_threshold):
        """Parameters of units small enough to stay gathered for the whole run (never re-fetched)."""
        return self.persistent_parameters, sum(p.numel() for p in self.persistent_parameters)

    def setup_zero_stage3_hooks(self):
        """Hooks are installed when the store is built; kept for call-site compatibility."""

    def pre_sub_module_forward_function(self, sub_module):
        self.get_param_coordinator(sub_module.training).fetch_sub_module(sub_module, forward=True)

    def post_sub_module_forward_function(self, sub_module):
        self.get_param_coordinator(sub_module.training).release_sub_module(sub_module, backward=False)

    def pre_sub_module_backward_function(self, sub_module):
        self.get_param_coordinator(True).fetch_sub_module(sub_module, forward=False)

    def post_sub_module_backward_function(self, sub_module):
        self.get_param_coordinator(True).release_sub_module(sub_module, backward=True)

    def _remove_module_hooks(self):
        for h in self._store._module_hooks:
            h.remove()
        self._store._module_hooks.clear()

    def destroy(self):
        self._remove_module_hooks()
