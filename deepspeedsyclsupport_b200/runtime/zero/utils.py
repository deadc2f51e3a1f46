"""ZeRO helper functions (parity: deepspeed/runtime/zero/utils.py:20-93)."""
import os
from typing import List

import torch

from ... import comm as dist
from ...utils.logging import logger


def _initialize_parameter_parallel_groups(parameter_parallel_size=None):
    """Split the data-parallel world into consecutive groups of `parameter_parallel_size` ranks; returns this rank's group."""
    data_parallel_size = int(dist.get_world_size())
    parameter_parallel_size = parameter_parallel_size or data_parallel_size
    logger.info(f"data_parallel_size: {data_parallel_size}, parameter_parallel_size: {parameter_parallel_size}")
    assert data_parallel_size % parameter_parallel_size == 0, "world size should be divisible by parameter parallel size"
    rank = dist.get_rank()
    my_group = None
    for i in range(data_parallel_size // parameter_parallel_size):
        ranks = range(i * parameter_parallel_size, (i + 1) * parameter_parallel_size)
        group = dist.new_group(ranks)
        if rank in ranks:
            my_group = group
    return my_group


class ZeRORuntimeException(Exception):
    pass


def _supported_optimizers():
    from ...ops.adagrad import DeepSpeedCPUAdagrad
    from ...ops.adam import DeepSpeedCPUAdam, FusedAdam
    from ...ops.lion import DeepSpeedCPULion, FusedLion
    return [torch.optim.Adam, torch.optim.AdamW, FusedAdam, DeepSpeedCPUAdam, torch.optim.Adagrad, DeepSpeedCPUAdagrad, DeepSpeedCPULion,
            FusedLion]


ZERO_SUPPORTED_OPTIMIZERS = _supported_optimizers()
try:  # apex is optional
    import apex  # noqa: F401
    if hasattr(apex, "optimizers") and hasattr(apex.optimizers, "FusedAdam"):
        ZERO_SUPPORTED_OPTIMIZERS.append(apex.optimizers.FusedAdam)
except ImportError:
    pass


def is_zero_supported_optimizer(optimizer):
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"Checking ZeRO support for optimizer={optimizer.__class__.__name__} type={type(optimizer)}")
    return type(optimizer) in ZERO_SUPPORTED_OPTIMIZERS


def get_lst_from_rank0(lst: List[int]) -> List[int]:
    """The list as rank 0 has it (every rank must pass a list of the same length)."""
    device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor(lst if dist.get_rank() == 0 else [-1] * len(lst), dtype=torch.int64, device=device)
    dist.broadcast(t, src=0, async_op=False)
    return t.cpu().tolist()


def assert_ints_same_as_other_ranks(ints: List[int]) -> None:
    """Debug aid for 'every rank must fetch parameters in the same order': raises if this rank's list differs from rank 0's."""
    rank0_ints = get_lst_from_rank0(ints)
    if ints != rank0_ints:
        raise RuntimeError(f"disagreement between rank0 and rank{dist.get_rank()}: rank0: {rank0_ints}, rank{dist.get_rank()}: {ints}")
