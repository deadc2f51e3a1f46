"""`zero.Init`, `zero.GatheredParameters`, external-parameter registration.

Parity: deepspeed/runtime/zero/partition_parameters.py:303-554 (module __init__ patching), :734-967 (Init),
:1998-2134 (GatheredParameters), :77-120 (register_external_parameter).

`Init` shards every parameter the moment its owning sub-module finishes constructing, so a model larger than one GPU
can be built.  At that point the block structure is not known yet, so parameters are first sharded individually
(`ceil(numel / world)` per rank, the reference's own layout); when the engine later builds the flat unit store it
re-assembles one unit at a time from these shards (sharded_optimizer._unit_flat_from_params), never materialising
more than one unit.
"""
import functools
import math

import torch

from ... import comm as dist
from ...accelerator import get_accelerator
from ...utils.logging import logger
from .param_store import ZeroParamStatus, ZeroParamType  # noqa: F401

_zero_init_context = 0
_orig_module_apply = None
_next_init_id = 0  # provisional ds_id given at construction (client code tests `hasattr(p, "ds_id")`); the engine renumbers


def is_zero_param(parameter):
    if not torch.is_tensor(parameter):
        return False
    return hasattr(parameter, "ds_id") or getattr(parameter, "ds_init_sharded", False)


def gather_init_sharded(p):
    """Full tensor of a parameter that was sharded by `Init` (one all-gather)."""
    world = p.ds_init_world
    shard = p.ds_tensor
    if world > 1:
        full = torch.empty(shard.numel() * world, dtype=shard.dtype, device=shard.device)
        dist.all_gather_into_tensor(full, shard, group=p.ds_process_group)
    else:
        full = shard
    return full[:p.ds_numel].view(p.ds_shape)


def _shard_one(p, group, world, rank, device, dtype):
    data = p.data
    if data.is_floating_point() and dtype is not None:
        data = data.to(dtype)
    data = data.to(device)
    if dist.is_initialized() and world > 1:
        dist.broadcast(data, src=dist.get_global_rank(group, 0), group=group)
    p.ds_shape = data.shape
    p.ds_numel = data.numel()
    per = (p.ds_numel + world - 1) // world
    flat = torch.zeros(per * world, dtype=data.dtype, device=device)
    flat[:p.ds_numel].copy_(data.reshape(-1))
    p.ds_tensor = flat[rank * per:(rank + 1) * per].clone()
    global _next_init_id
    p.ds_id = _next_init_id
    _next_init_id += 1
    p.ds_init_sharded = True
    p.ds_init_world = world
    p.ds_process_group = group
    p.ds_status = ZeroParamStatus.NOT_AVAILABLE
    p.ds_persist = False
    p.data = torch.empty(0, dtype=data.dtype, device=device)


class Init:
    """Context manager: parameters of modules constructed inside are sharded across the data-parallel group at
    construction time.  Arguments follow the reference (`deepspeed.zero.Init`)."""

    def __init__(self, module=None, data_parallel_group=None, mem_efficient_linear=True, remote_device=None, pin_memory=False,
                 config_dict_or_path=None, config=None, enabled=True, dtype=None, mpu=None, zero_param_parallel_group=None,
                 zero_quantized_weights=False, zero_quantized_nontrainable_weights=False, sequence_data_parallel_group=None,
                 param_swapper=None):
        self.enabled = enabled
        self.dtype = dtype
        cfg = config_dict_or_path if config_dict_or_path is not None else config
        if cfg is not None and dtype is None:
            from ..config import DeepSpeedConfig, _load_config_dict
            d = _load_config_dict(cfg)
            if d.get("bf16", {}).get("enabled", False):
                self.dtype = torch.bfloat16
            elif d.get("fp16", {}).get("enabled", False):
                self.dtype = torch.float16
        if enabled and not dist.is_initialized():
            dist.init_distributed()
        self.group = sequence_data_parallel_group or data_parallel_group or (dist.get_world_group() if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        self.device = torch.device(get_accelerator().current_device_name())
        self.quantized_weights = zero_quantized_weights
        self._patched = []
        if module is not None and enabled:
            self._convert_module(module)

    # -- conversion
    def _convert_module(self, module):
        for m in module.modules():
            self._post_init(m)

    def _post_init(self, module):
        for name, p in module.named_parameters(recurse=False):
            if not is_zero_param(p):
                _shard_one(p, self.group, self.world, self.rank, self.device, self.dtype)

    # -- group introspection (reference: partition_parameters.py:1981-1995)
    def get_partition_dp_group(self, param):
        return getattr(param, "ds_process_group", self.group)

    def get_partition_rank(self):
        """Relative rank inside the parameter-partition group (MiCS overloads the group)."""
        return self.rank

    @property
    def num_partitions(self):
        return self.world

    def get_dp_process_group(self):
        """The communication group with all data-parallel ranks."""
        return self.group

    # -- nn.Module patching
    def __enter__(self):
        global _zero_init_context, _orig_module_apply
        if not self.enabled:
            return self
        _zero_init_context += 1
        init_ctx = self

        def wrap_init(cls):
            if "__init__" not in cls.__dict__ or getattr(cls.__init__, "_ds_wrapped", False):
                return
            orig = cls.__init__

            @functools.wraps(orig)
            def wrapper(module, *args, **kwargs):
                module._ds_init_depth = getattr(module, "_ds_init_depth", 0) + 1
                try:
                    orig(module, *args, **kwargs)
                finally:
                    module._ds_init_depth -= 1
                if module._ds_init_depth == 0 and _zero_init_context > 0:
                    init_ctx._post_init(module)

            wrapper._ds_wrapped = True
            wrapper._ds_orig = orig
            cls.__init__ = wrapper
            init_ctx._patched.append(cls)

        def all_subclasses(cls):
            seen, stack = set(), [cls]
            while stack:
                c = stack.pop()
                for s in c.__subclasses__():
                    if s not in seen:
                        seen.add(s)
                        stack.append(s)
            return seen

        for cls in all_subclasses(torch.nn.Module):
            wrap_init(cls)
        self._orig_init_subclass = torch.nn.Module.__dict__.get("__init_subclass__")

        def _init_subclass(cls, **kwargs):
            wrap_init(cls)

        torch.nn.Module.__init_subclass__ = classmethod(_init_subclass)
        # module.apply(fn) must see full parameters (weight-init functions), reference: override_module_apply
        if _orig_module_apply is None:
            _orig_module_apply = torch.nn.Module.apply

            def apply_with_gather(module, fn):
                for child in module.children():
                    child.apply(fn)
                params = [p for p in module.parameters(recurse=False) if getattr(p, "ds_init_sharded", False)]
                if params:
                    with GatheredParameters(params, modifier_rank=0):
                        fn(module)
                else:
                    fn(module)
                return module

            torch.nn.Module.apply = apply_with_gather
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        global _zero_init_context, _orig_module_apply
        if not self.enabled:
            return False
        _zero_init_context -= 1
        for cls in self._patched:
            if getattr(cls.__init__, "_ds_wrapped", False):
                cls.__init__ = cls.__init__._ds_orig
        self._patched = []
        if self._orig_init_subclass is not None:
            torch.nn.Module.__init_subclass__ = self._orig_init_subclass
        else:
            try:
                del torch.nn.Module.__init_subclass__
            except AttributeError:
                pass
        if _zero_init_context == 0 and _orig_module_apply is not None:
            torch.nn.Module.apply = _orig_module_apply
            _orig_module_apply = None
        return False


def shutdown_init_context():
    """`initialize()` suspends an enclosing Init while it builds the engine (parity: partition_parameters.py:530-541)."""
    global _zero_init_context
    _saved.append(_zero_init_context)
    _zero_init_context = 0


def restore_init_context():
    global _zero_init_context
    if _saved:
        _zero_init_context = _saved.pop()


_saved = []


class GatheredParameters:
    """Materialise full parameters inside the context.  If `modifier_rank` is given, that rank's modifications are
    broadcast and written back to the shards on exit."""

    def __init__(self, params, modifier_rank=None, fwd_module=None, enabled=True):
        self.enabled = enabled
        if not enabled:
            return
        if isinstance(params, torch.nn.Module):
            params = list(params.parameters())
        elif torch.is_tensor(params):
            params = [params]
        else:
            params = list(params)
        self.params = [p for p in params if is_zero_param(p)]
        self.modifier_rank = modifier_rank
        self.fwd_module = fwd_module
        if fwd_module is not None:
            for p in self.params:
                register_external_parameter(fwd_module, p)

    def __enter__(self):
        if not self.enabled:
            return
        self.init_params = [p for p in self.params if getattr(p, "ds_init_sharded", False)]
        self.unit_params = [p for p in self.params if not getattr(p, "ds_init_sharded", False) and hasattr(p, "ds_unit")]
        for p in self.init_params:
            p.data = gather_init_sharded(p).clone()
            p.ds_status = ZeroParamStatus.AVAILABLE
        self._units = {}
        for p in self.unit_params:
            u = p.ds_unit
            if id(u) not in self._units:
                self._units[id(u)] = (u, u.status)
                p.ds_zero_optimizer._fetch(u)

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        for p in self.init_params:
            if self.modifier_rank is not None:
                if dist.is_initialized() and p.ds_init_world > 1:
                    dist.broadcast(p.data, src=dist.get_global_rank(p.ds_process_group, self.modifier_rank), group=p.ds_process_group)
                per = p.ds_tensor.numel()
                rank = dist.get_rank(p.ds_process_group) if dist.is_initialized() else 0
                flat = torch.zeros(per * p.ds_init_world, dtype=p.data.dtype, device=p.data.device)
                flat[:p.ds_numel].copy_(p.data.reshape(-1))
                p.ds_tensor.copy_(flat[rank * per:(rank + 1) * per])
            p.data = torch.empty(0, dtype=p.data.dtype, device=p.data.device)
            p.ds_status = ZeroParamStatus.NOT_AVAILABLE
        for u, was in self._units.values():
            opt = u.params[0].ds_zero_optimizer
            if self.modifier_rank is not None:
                if dist.is_initialized() and opt.shard_world > 1:
                    dist.broadcast(u.full, src=dist.get_global_rank(opt.shard_group, self.modifier_rank), group=opt.shard_group)
                opt._write_back_unit(u)
            if was == ZeroParamStatus.NOT_AVAILABLE:
                opt.release_unit(u)
        return False


def register_external_parameter(module, parameter):
    """Declare that `module.forward` uses `parameter` although another module owns it (tied weights etc.)."""
    if not isinstance(parameter, torch.nn.Parameter):
        raise RuntimeError("Parameter is not a torch.nn.Parameter")
    if not hasattr(module, "_external_params"):
        module._external_params = {}
    module._external_params[id(parameter)] = parameter
    opt = getattr(parameter, "ds_zero_optimizer", None)
    if opt is not None and opt.stage == 3 and opt.shard_world > 1:
        opt._hook_module(module, [parameter.ds_unit])


def unregister_external_parameter(module, parameter):
    if hasattr(module, "_external_params"):
        module._external_params.pop(id(parameter), None)
