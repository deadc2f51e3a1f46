"""Launcher-environment and call-introspection helpers of the comm layer (parity: deepspeed/comm/utils.py:11-134)."""
import inspect
import os

from ..utils.logging import logger


def _from_env(names, what, default):
    for n in names:
        v = os.environ.get(n)
        if v is not None:
            return int(v)
    logger.debug(f"{what} not set by a launcher ({', '.join(names)}); assuming {default}")
    return default


def get_local_rank_from_launcher():
    # DeepSpeed / torchrun launchers, then Open MPI, then Slurm
    return _from_env(("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID"), "local rank", 0)


def get_world_rank_from_launcher():
    return _from_env(("RANK", "OMPI_COMM_WORLD_RANK", "SLURM_PROCID"), "rank", 0)


def get_world_size_from_launcher():
    return _from_env(("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "SLURM_NPROCS"), "world size", 1)


def get_default_args(func):
    return {k: v.default for k, v in inspect.signature(func).parameters.items() if v.default is not inspect.Parameter.empty}


def get_tensor_position(func):
    """Index of the argument that carries the payload of a collective (by the names the comm API uses)."""
    names = list(inspect.signature(func).parameters)
    for candidate in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensor"):
        if candidate in names:
            return names.index(candidate)
    return -1


def get_tensor_kwarg(func, kwargs):
    merged = dict(get_default_args(func), **kwargs)
    for candidate in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensor"):
        if merged.get(candidate) is not None:
            return merged[candidate]
    return None


def get_msg_size_from_args(func, *args, **kwargs):
    """Bytes moved by one call of a collective, however the payload was passed (positional or keyword, tensor or list)."""
    import torch
    pos = get_tensor_position(func)
    payload = args[pos] if 0 <= pos < len(args) else get_tensor_kwarg(func, kwargs)
    if payload is None:
        return 0
    if isinstance(payload, torch.Tensor):
        return payload.element_size() * payload.nelement()
    return sum(t.element_size() * t.nelement() for t in payload)


def get_debug_log_name(func_args, debug):
    return func_args["log_name"] + " | [Caller Func: " + debug + "]" if debug else func_args["log_name"]
