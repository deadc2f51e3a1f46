"""Access to the full (un-sharded) fp32 weights / gradients / optimizer state of a parameter under any ZeRO stage.
Parity: deepspeed/utils/tensor_fragment.py:101-255 (safe_get/set_full_*, safe_get/set_local_*)."""
import torch


def _opt(param):
    return getattr(param, "ds_zero_optimizer", None)


def safe_get_full_fp32_param(param):
    opt = _opt(param)
    if opt is None:
        return param.data.float()
    return opt.get_full_fp32_param(param)


def safe_set_full_fp32_param(param, value):
    opt = _opt(param)
    if opt is None:
        param.data.copy_(value)
    else:
        opt.set_full_fp32_param(param, value)


def safe_get_full_optimizer_state(param, optim_state_key):
    opt = _opt(param)
    if opt is None:
        return None
    return opt.get_full_optimizer_state(param, optim_state_key)


def safe_set_full_optimizer_state(param, value, optim_state_key):
    opt = _opt(param)
    if opt is not None:
        opt.set_full_optimizer_state(param, value, optim_state_key)


def safe_get_full_grad(param):
    opt = _opt(param)
    if opt is None:
        return param.grad
    return opt.get_full_grad(param)


def safe_get_local_fp32_param(param):
    opt = _opt(param)
    u = param.ds_unit
    off, n = u.param_slice(param.ds_index)
    lo, hi = u.shard_range()
    a, b = max(off, lo), min(off + n, hi)
    return opt._master_slice(u.shard_off + a - lo, u.shard_off + b - lo) if a < b else torch.empty(0)


def safe_get_local_grad(param):
    opt = _opt(param)
    u = param.ds_unit
    off, n = u.param_slice(param.ds_index)
    lo, hi = u.shard_range()
    a, b = max(off, lo), min(off + n, hi)
    return opt.grad_shard[u.shard_off + a - lo:u.shard_off + b - lo].float() if a < b else torch.empty(0)


def safe_get_local_optimizer_state(param, optim_state_key):
    opt = _opt(param)
    u = param.ds_unit
    off, n = u.param_slice(param.ds_index)
    lo, hi = u.shard_range()
    a, b = max(off, lo), min(off + n, hi)
    return opt._state_slice(optim_state_key, u.shard_off + a - lo, u.shard_off + b - lo) if a < b else torch.empty(0)


def safe_set_local_fp32_param(param, value):
    opt = _opt(param)
    u = param.ds_unit
    off, n = u.param_slice(param.ds_index)
    lo, hi = u.shard_range()
    a, b = max(off, lo), min(off + n, hi)
    if a < b:
        s, e = u.shard_off + a - lo, u.shard_off + b - lo
        for r in opt.regions:
            x, y = max(s, r.lo), min(e, r.hi)
            if x < y:
                r.master[x - r.lo:y - r.lo].copy_(value.reshape(-1)[x - s:y - s])
        opt.lp_shard[s:e].copy_(value.reshape(-1))


def safe_set_local_optimizer_state(param, value, optim_state_key):
    opt = _opt(param)
    u = param.ds_unit
    off, n = u.param_slice(param.ds_index)
    lo, hi = u.shard_range()
    a, b = max(off, lo), min(off + n, hi)
    if a < b:
        s, e = u.shard_off + a - lo, u.shard_off + b - lo
        for r in opt.regions:
            opt._ensure_state(r, opt._opt_kind)
            x, y = max(s, r.lo), min(e, r.hi)
            if x < y:
                r.state[optim_state_key][x - r.lo:y - r.lo].copy_(value.reshape(-1)[x - s:y - s])


# ---------------------------------------------------------------------------------------------------------------------
# Fragment bookkeeping types (parity: utils/tensor_fragment.py:14-98, 258-296).  The flat unit store resolves
# offsets itself, but checkpoint converters and user code describe "which slice of a parameter does this rank's
# partition hold" with these records, so they are provided with the same fields.
# ---------------------------------------------------------------------------------------------------------------------
from dataclasses import dataclass, field
from typing import Dict, Optional


@dataclass
class fragment_address:
    numel: int
    start: int


@dataclass
class tensor_fragment:
    lp_fragment: torch.Tensor
    lp_fragment_address: fragment_address
    hp_fragment: torch.Tensor
    hp_fragment_address: fragment_address
    optim_fragment: Dict = field(default_factory=dict)
    gradient_dict: Optional[Dict] = None
    offload_gradient_dict: Optional[Dict] = None
    use_offload: bool = False
    param_group_index: int = 0

    def update_hp(self):
        self.hp_fragment.data.copy_(self.lp_fragment.data)

    def update_lp(self):
        self.lp_fragment.data.copy_(self.hp_fragment.data)

    def get_optim_state_fragment(self, key):
        if key not in self.optim_fragment:
            raise ValueError(f"{key} not found in optimizer state fragment")
        return self.optim_fragment[key]

    def get_hp_fragment_address(self):
        return self.hp_fragment_address

    def get_optim_state_keys(self):
        return list(self.optim_fragment.keys())

    def get_hp_fragment(self, optim_state_key=None):
        return self.hp_fragment if optim_state_key is None else self.get_optim_state_fragment(optim_state_key)


def get_hp_fragment_mapping(lp_param, lp_start, flat_hp_partition, gradient_dict, offload_gradient_dict, use_offload,
                            param_group_index, partition_start, partition_size, optimizer_state_dict=None):
    """Describe the overlap of ``lp_param`` (occupying ``[lp_start, lp_start + numel)`` of the flat group) with this
    rank's fp32 partition ``[partition_start, partition_start + partition_size)``: views of both sides plus the same
    slice of every optimizer-state tensor shaped like the partition."""
    lp_end = lp_start + lp_param.numel()
    lo, hi = max(lp_start, partition_start), min(lp_end, partition_start + partition_size)
    assert lo < hi, f"no overlap: lp [{lp_start}, {lp_end}) vs partition [{partition_start}, {partition_start + partition_size})"
    n = hi - lo
    hp_addr = fragment_address(numel=n, start=lo - partition_start)
    lp_addr = fragment_address(numel=n, start=lo - lp_start)
    hp_frag = flat_hp_partition.narrow(0, hp_addr.start, n)
    lp_frag = lp_param.data.flatten().narrow(0, lp_addr.start, n) if lp_param.is_contiguous() \
        else lp_param.data.reshape(-1).narrow(0, lp_addr.start, n)
    optim = {}
    for key, value in (optimizer_state_dict or {}).items():
        if torch.is_tensor(value) and value.dim() > 0 and value.numel() == flat_hp_partition.numel():
            optim[key] = value.narrow(0, hp_addr.start, n)
    return tensor_fragment(lp_fragment=lp_frag, lp_fragment_address=lp_addr, hp_fragment=hp_frag,
                           hp_fragment_address=hp_addr, optim_fragment=optim, gradient_dict=gradient_dict,
                           offload_gradient_dict=offload_gradient_dict, use_offload=use_offload,
                           param_group_index=param_group_index)


def get_full_hp_param(param, optim_state_key=None):
    """Function form of the method the reference binds onto low-precision parameters."""
    if optim_state_key is None:
        return safe_get_full_fp32_param(param)
    return safe_get_full_optimizer_state(param, optim_state_key)


def set_full_hp_param(param, value, optim_state_key=None):
    if optim_state_key is None:
        safe_set_full_fp32_param(param, value)
    else:
        safe_set_full_optimizer_state(param, value, optim_state_key)


def get_full_hp_grad(param):
    return safe_get_full_grad(param)
