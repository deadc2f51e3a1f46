"""MoE helpers: parameter classification, param-group splitting, the ZeRO composite optimizer for models with
experts, and expert checkpoint files.  Parity: deepspeed/moe/utils.py, engine.py:3151-3281 (_save_moe_checkpoint),
engine.py:2417-2433 / stage_1_and_2.py:1035-1038 (expert gradients reduced over the expert-data-parallel group)."""
import os
import re
from collections import OrderedDict

import torch

from .. import comm as dist
from ..utils import groups
from ..utils.logging import logger


def has_moe_layers(m):
    from .layer import MoE
    has_moe, num_experts = False, 0
    for _, module in m.named_modules():
        if isinstance(module, MoE):
            has_moe = True
            num_experts = module.num_experts
            break
    return has_moe, num_experts


def is_moe_param(param: torch.Tensor) -> bool:
    return hasattr(param, "allreduce") and not param.allreduce


def split_params_into_shared_and_expert_params(params):
    shared_params, expert_params = [], []
    for p in params:
        (expert_params if is_moe_param(p) else shared_params).append(p)
    return shared_params, expert_params


def split_params_grads_into_shared_and_expert_params(group):
    expert_grads, shared_grads = [], []
    for p in group:
        if p.grad is not None:
            (expert_grads if is_moe_param(p) else shared_grads).append(p.grad.to(p.dtype))
    return shared_grads, expert_grads


def split_params_into_different_moe_groups_for_optimizer(param_groups, max_group_size=178956971):
    """Split every optimizer param group into a dense group + one group per expert-parallel group name."""
    if isinstance(param_groups, tuple):
        param_groups = list(param_groups)
    elif isinstance(param_groups, dict):
        param_groups = [param_groups]
    elif not isinstance(param_groups, list):
        raise ValueError(f"Unknown param group type of {type(param_groups)}")
    data_parallel_group_names = set()
    for param_group in param_groups:
        for param in param_group["params"]:
            if is_moe_param(param):
                data_parallel_group_names.add(param.group_name)
    data_parallel_group_names = sorted(data_parallel_group_names)
    group_moe = {}
    for param_group in param_groups:
        group_moe[param_group["name"]] = {}
        for key in data_parallel_group_names:
            group_moe[param_group["name"]][key] = {"name": key, "moe": True, "params": []}
            for ori_key in param_group.keys():
                if ori_key not in ("name", "params"):
                    group_moe[param_group["name"]][key][ori_key] = param_group[ori_key]
    for param_group in param_groups:
        new_params = []
        for param in param_group["params"]:
            if is_moe_param(param):
                group_moe[param_group["name"]][param.group_name]["params"].append(param)
            else:
                new_params.append(param)
        param_group["params"] = new_params
    for k, v in group_moe.items():
        for k1, v1 in v.items():
            cur_group, all_groups, size_of_cur_group = [], [], 0
            for param in v1["params"]:
                if size_of_cur_group + param.numel() <= max_group_size:
                    cur_group.append(param)
                    size_of_cur_group += param.numel()
                else:
                    all_groups.append(cur_group)
                    cur_group = [param]
                    size_of_cur_group = param.numel()
            if cur_group:
                all_groups.append(cur_group)
            for group in all_groups:
                new_dict = {kk: vv for kk, vv in v1.items() if kk != "params"}
                new_dict["params"] = group
                param_groups.append(new_dict)
    return tuple(param_groups)


class MoEZeroOptimizer:
    """ZeRO for models with experts: dense parameters are sharded / reduced over the data-parallel group, expert
    parameters over their expert-data-parallel group; both clip against ONE global gradient norm."""

    def __init__(self, module, init_optimizer, *, stage, dp_group, expert_data_parallel_group, **common):
        from ..runtime.zero.sharded_optimizer import ZeroShardedOptimizer
        assert stage <= 2, "MoE is not supported with ZeRO stage 3 (reference: engine.py:1535)"
        kw = dict(common)
        kw["grad_accum_dtype"] = kw.pop("gradient_accumulation_dtype", None)
        post = kw.pop("postscale_gradients", True)
        kw["prescale_gradients"] = not post
        self.dense = ZeroShardedOptimizer(module, init_optimizer, stage=stage, dp_group=dp_group,
                                          param_filter=lambda p: not is_moe_param(p), **kw)
        self.expert = {}
        names = sorted({p.group_name for p in module.parameters() if is_moe_param(p)})
        for name in names:
            self.expert[name] = ZeroShardedOptimizer(module, init_optimizer, stage=stage, dp_group=expert_data_parallel_group[name],
                                                     param_filter=lambda p, n=name: is_moe_param(p) and p.group_name == n, **kw)
        self.optimizer = init_optimizer
        self.param_groups = init_optimizer.param_groups
        self.stage = stage
        self.partition_count = self.dense.partition_count
        self.dp_group = dp_group
        self.shard_world = self.dense.shard_world
        self.overflow = False
        self.custom_loss_scaler = False

    def _all(self):
        return [self.dense] + list(self.expert.values())

    @property
    def units(self):
        return [u for o in self._all() for u in o.units]

    @property
    def loss_scale(self):
        return self.dense.loss_scale

    cur_scale = loss_scale

    def backward(self, loss, retain_graph=False):
        self.dense.backward(loss, retain_graph=retain_graph)

    def end_micro_step(self, boundary, local_only=False):
        for o in self._all():
            o.end_micro_step(boundary, local_only=local_only)

    def zero_grad(self, set_to_none=True):
        for o in self._all():
            o.zero_grad()

    def step(self, closure=None):
        skips = [o.begin_step() for o in self._all()]
        self.overflow = any(skips)
        if self.overflow:
            for o, s in zip(self._all(), skips):
                if not s:
                    o.zero_grad()
            return
        sq = self.dense._compute_global_norm().square()
        for name, o in self.expert.items():
            e = o._compute_global_norm().square()
            ep_group = groups._get_expert_parallel_group(name)
            if dist.get_world_size(ep_group) > 1:
                dist.all_reduce(e, group=ep_group)  # experts differ across expert-parallel ranks
            sq = sq + e
        norm = sq.sqrt()
        for o in self._all():
            o.finish_step(norm)

    def get_global_grad_norm(self):
        return self.dense.get_global_grad_norm()

    def state_dict(self):
        return {"dense": self.dense.state_dict(), "expert": {k: v.state_dict() for k, v in self.expert.items()},
                "partition_count": self.partition_count, "zero_stage": self.stage}

    def load_state_dict(self, sd, **kw):
        if isinstance(sd, (list, tuple)):
            sd = sd[0]
        self.dense.load_state_dict(sd["dense"], **kw)
        for k, v in self.expert.items():
            v.load_state_dict(sd["expert"][k], **kw)

    def refresh_from_params(self):
        for o in self._all():
            o.refresh_from_params()

    def checkpoint_event_prologue(self):
        pass

    def checkpoint_event_epilogue(self):
        pass

    def get_full_fp32_param(self, p):
        return p.ds_zero_optimizer.get_full_fp32_param(p)


def _expert_key_pattern():
    return re.compile(r"(.*deepspeed_moe\.experts\.deepspeed_experts\.)(\d+)(\..*)")


def save_moe_checkpoint(engine, save_dir, tag, client_state, exclude_frozen_parameters=False):
    """Per-expert files layer_{L}_expert_{E}_mp_rank_XX_model_states.pt (global expert ids) + the dense model states
    on dp rank 0 + expp_rank_{R}_mp_rank_XX_optim_states.pt per expert-parallel rank (reference layout)."""
    from .layer import MoE
    full_sd = engine.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
    moe_layer_id = 0
    expert_keys = set()
    for n_module, module in engine.module.named_modules():
        if not isinstance(module, MoE):
            continue
        group_name = module.expert_group_name
        num_local_experts = module.num_local_experts
        expp_rank = groups._get_expert_parallel_rank(group_name)
        exp_dp_rank = groups._get_expert_data_parallel_rank(group_name)
        prefix = f"{n_module}.deepspeed_moe.experts.deepspeed_experts."
        if exp_dp_rank == 0:
            for local_expert_id in range(num_local_experts):
                global_expert_id = expp_rank * num_local_experts + local_expert_id
                expert_sd = OrderedDict()
                for key, val in full_sd.items():
                    lp = f"{prefix}{local_expert_id}."
                    if key.startswith(lp):
                        expert_sd[f"{prefix}{global_expert_id}." + key[len(lp):]] = val.clone().detach()
                path = engine._get_expert_ckpt_name(save_dir, moe_layer_id, global_expert_id, tag, 0)
                engine.checkpoint_engine.save(expert_sd, path)
        for key in full_sd.keys():
            if key.startswith(prefix):
                expert_keys.add(key)
        moe_layer_id += 1
    largest_group_name = groups._get_max_expert_size_name()
    expp_rank = groups._get_expert_parallel_rank(largest_group_name)
    exp_dp_rank = groups._get_expert_data_parallel_rank(largest_group_name)
    if exp_dp_rank == 0 and engine.optimizer is not None:
        optimizer_state = {"optimizer": engine.optimizer.state_dict()}
        engine.checkpoint_engine.save(optimizer_state, engine._get_optimizer_ckpt_name(save_dir, tag, expp_rank))
    if groups._get_data_parallel_rank() == 0:
        dense_sd = OrderedDict((k, v) for k, v in full_sd.items() if k not in expert_keys)
        state = {"module": dense_sd, "lr_scheduler": engine.lr_scheduler.state_dict() if engine.lr_scheduler is not None else None,
                 "sparse_tensor_module_names": [], "skipped_steps": engine.skipped_steps, "global_steps": engine.global_steps,
                 "global_samples": engine.global_samples, "dp_world_size": engine.dp_world_size, "mp_world_size": engine.mp_world_size,
                 "num_experts": engine.num_experts, "ds_config": engine._config._param_dict}
        state.update(client_state)
        engine.checkpoint_engine.save(state, engine._get_ckpt_name(save_dir, tag))


def load_moe_state_dict(engine, load_dir, tag, checkpoint):
    """Merge the per-expert files of this rank's experts back into checkpoint['module'] (local expert ids)."""
    from .layer import MoE
    moe_layer_id = 0
    sd = checkpoint["module"]
    for n_module, module in engine.module.named_modules():
        if not isinstance(module, MoE):
            continue
        group_name = module.expert_group_name
        num_local_experts = module.num_local_experts
        expp_rank = groups._get_expert_parallel_rank(group_name)
        prefix = f"{n_module}.deepspeed_moe.experts.deepspeed_experts."
        for local_expert_id in range(num_local_experts):
            global_expert_id = expp_rank * num_local_experts + local_expert_id
            path = engine._get_expert_ckpt_name(load_dir, moe_layer_id, global_expert_id, tag, 0)
            expert_sd = engine.checkpoint_engine.load(path, map_location="cpu")
            gp = f"{prefix}{global_expert_id}."
            for key, val in expert_sd.items():
                if key.startswith(gp):
                    sd[f"{prefix}{local_expert_id}." + key[len(gp):]] = val
        moe_layer_id += 1
    opt_path = engine._get_optimizer_ckpt_name(load_dir, tag, groups._get_expert_parallel_rank(groups._get_max_expert_size_name()))
    if os.path.exists(opt_path) and engine.optimizer is not None:
        engine._moe_optimizer_state = engine.checkpoint_engine.load(opt_path, map_location="cpu")["optimizer"]


def load_moe_expert_files(checkpoint_path, tag, state_dict, old_moe_load, model, mpu=None, num_experts=1, checkpoint_engine=None):
    """Engine-free form of the expert merge (reference: DeepSpeedEngine.load_moe_state_dict, a static method used by the
    inference engine as well): reads `layer_{L}_expert_{E}_mp_rank_{R}_model_states.pt` for this rank's experts and writes
    their tensors into `state_dict` under local expert ids.  `old_moe_load`: one expert-parallel group with `num_experts[0]`
    experts and file names without the layer id grouping of newer checkpoints."""
    from .layer import MoE
    if checkpoint_engine is None:
        from ..runtime.checkpoint_engine.torch_checkpoint_engine import TorchCheckpointEngine
        checkpoint_engine = TorchCheckpointEngine()
    mp_rank = 0 if mpu is None else mpu.get_model_parallel_rank()

    def expert_file(layer_id, expert_id):
        name = f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt"
        return os.path.join(checkpoint_path, name if tag is None else os.path.join(str(tag), name))

    if old_moe_load:
        expp_rank = groups._get_expert_data_parallel_rank(groups._get_max_expert_size_name())
        n = num_experts[0] if isinstance(num_experts, (list, tuple)) else num_experts
        num_local = max(n // groups._get_expert_parallel_world_size(groups._get_max_expert_size_name()), 1)
        for local in range(num_local):
            glob = expp_rank * num_local + local
            expert_sd = checkpoint_engine.load(expert_file(0, glob), map_location="cpu")
            for key in list(expert_sd.keys()):
                state_dict[key.replace(f"expert_{glob}", f"deepspeed_moe.experts.deepspeed_experts.{local}")] = expert_sd.pop(key)
        return state_dict
    layer_id = 0
    for n_module, module in model.named_modules():
        if not isinstance(module, MoE):
            continue
        expp_rank = groups._get_expert_parallel_rank(module.expert_group_name)
        prefix = f"{n_module}.deepspeed_moe.experts.deepspeed_experts."
        for local in range(module.num_local_experts):
            glob = expp_rank * module.num_local_experts + local
            expert_sd = checkpoint_engine.load(expert_file(layer_id, glob), map_location="cpu")
            gp = f"{prefix}{glob}."
            for key, val in expert_sd.items():
                if key.startswith(gp):
                    state_dict[f"{prefix}{local}." + key[len(gp):]] = val
        layer_id += 1
    return state_dict
