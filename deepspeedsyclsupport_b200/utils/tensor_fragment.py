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
