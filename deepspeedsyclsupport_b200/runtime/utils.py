"""Misc runtime utilities: partitioning helpers, norms, clip, memory report, DummyOptim.
Parity: deepspeed/runtime/utils.py:39-1041."""
import gc
import math
import os
from bisect import bisect_left

import torch

from .. import comm as dist
from ..accelerator import get_accelerator
from ..utils.logging import logger


class DummyOptim:
    """Placeholder optimizer when ZeRO-3 is used without a real optimizer (inference)."""

    def __init__(self, params):
        self.param_groups = [{"params": params}]


def noop_decorator(func):
    return func


def ensure_directory_exists(filename):
    dirname = os.path.dirname(filename)
    os.makedirs(dirname, exist_ok=True)


def set_random_seed(seed):
    import random

    import numpy
    random.seed(seed)
    numpy.random.seed(seed)
    torch.manual_seed(seed)


def is_model_parallel_parameter(p) -> bool:
    if hasattr(p, "model_parallel") and p.model_parallel:
        return True
    if hasattr(p, "tensor_model_parallel") and p.tensor_model_parallel:
        return True
    return False


def bwc_tensor_model_parallel_rank(mpu=None):
    if mpu is None:
        return 0
    if hasattr(mpu, "get_tensor_model_parallel_rank"):
        return mpu.get_tensor_model_parallel_rank()
    if hasattr(mpu, "get_slice_parallel_rank"):
        return mpu.get_slice_parallel_rank()
    return mpu.get_model_parallel_rank()


def move_to_device(item, device, criterion_func=None):
    if torch.is_tensor(item) and (criterion_func is None or criterion_func(item)):
        return item.to(device)
    if isinstance(item, list):
        return [move_to_device(v, device, criterion_func) for v in item]
    if isinstance(item, tuple):
        return tuple(move_to_device(v, device, criterion_func) for v in item)
    if isinstance(item, dict):
        return {k: move_to_device(v, device, criterion_func) for k, v in item.items()}
    return item


def get_global_norm(norm_list):
    total_norm = 0.0
    for norm in norm_list:
        total_norm += norm ** 2.0
    return math.sqrt(total_norm)


def get_grad_norm(parameters, norm_type=2, mpu=None):
    """MP-aware gradient norm over `parameters` (duplicated tensors counted once: only mp rank 0 counts them)."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = [p for p in parameters if p.grad is not None]
    norm_type = float(norm_type)
    dev = get_accelerator().current_device_name()
    if norm_type == float("inf"):
        total_norm = max((p.grad.data.abs().max() for p in parameters), default=torch.zeros((), device=dev))
        total_norm_t = torch.as_tensor(float(total_norm), device=dev).reshape(1)
        if mpu is not None:
            dist.all_reduce(total_norm_t, op=dist.ReduceOp.MAX, group=mpu.get_model_parallel_group())
        total_norm = total_norm_t[0].item()
    else:
        total_norm = 0.0
        tmp_rank = bwc_tensor_model_parallel_rank(mpu=mpu)
        for p in parameters:
            if hasattr(p, "ds_pipe_replicated") and p.ds_pipe_replicated:
                continue
            if (tmp_rank > 0) and not is_model_parallel_parameter(p):
                continue
            total_norm += p.grad.data.float().norm(norm_type).item() ** norm_type
        total_norm_t = torch.as_tensor(float(total_norm), device=dev).reshape(1)
        if mpu is not None:
            dist.all_reduce(total_norm_t, op=dist.ReduceOp.SUM, group=mpu.get_model_parallel_group())
        total_norm = total_norm_t[0].item() ** (1.0 / norm_type)
    if total_norm == float("inf") or total_norm == -float("inf") or total_norm != total_norm:
        total_norm = -1
    return total_norm


def get_weight_norm(parameters, norm_type=2, mpu=None):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    norm_type = float(norm_type)
    dev = get_accelerator().current_device_name()
    if norm_type == float("inf"):
        total_norm = max(float(p.data.abs().max()) for p in parameters)
    else:
        total_norm = 0.0
        tmp_rank = bwc_tensor_model_parallel_rank(mpu=mpu)
        for p in parameters:
            if (tmp_rank > 0) and not is_model_parallel_parameter(p):
                continue
            total_norm += p.data.float().norm(norm_type).item() ** norm_type
        t = torch.as_tensor(float(total_norm), device=dev).reshape(1)
        if mpu is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=mpu.get_model_parallel_group())
        total_norm = t[0].item() ** (1.0 / norm_type)
    if total_norm == float("inf") or total_norm == -float("inf") or total_norm != total_norm:
        total_norm = -1
    return total_norm


def clip_grad_norm_(parameters, max_norm, norm_type=2, mpu=None):
    """Clip by global norm (MP-aware). Returns the total norm."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = list(filter(lambda p: p.grad is not None, parameters))
    total_norm = get_grad_norm(parameters, norm_type, mpu)
    if total_norm < 0:
        return total_norm
    clip_coef = max_norm / (total_norm + 1e-6)
    if clip_coef < 1:
        for p in parameters:
            p.grad.data.mul_(clip_coef)
    return total_norm


def clip_gradients(parameters, max_norm=1.0, global_grad_norm=None, mpu=None, eps=1e-6):
    if global_grad_norm is None:
        global_grad_norm = get_grad_norm(parameters, mpu=mpu)
    clip_coef = max_norm / (global_grad_norm + eps)
    if clip_coef < 1:
        for p in parameters:
            p.grad.detach().mul_(clip_coef)
    return global_grad_norm


def prefix_sum_inc(weights):
    weights_ = [w for w in weights]
    for x in range(1, len(weights_)):
        weights_[x] += weights_[x - 1]
    return weights_


def partition_uniform(num_items, num_parts):
    parts = [0] * (num_parts + 1)
    if num_items <= num_parts:
        for p in range(num_parts + 1):
            parts[p] = min(p, num_items)
        return parts
    chunksize = num_items // num_parts
    residual = num_items - (chunksize * num_parts)
    parts = list(range(0, (num_parts + 1) * chunksize, chunksize))
    for i in range(residual):
        parts[i + 1:] = [p + 1 for p in parts[i + 1:]]
    return parts


def _lprobe(weights, num_parts, bottleneck):
    num_items = len(weights)
    total_weight = weights[-1]
    parts = [0] * (num_parts + 1)
    for p in range(1, num_parts + 1):
        parts[p] = num_items
    bsum = bottleneck
    chunksize = num_items // num_parts
    step = chunksize
    for p in range(1, num_parts):
        while (step < num_items) and (weights[step] < bsum):
            step += chunksize
        parts[p] = bisect_left(weights, bsum, lo=step - chunksize, hi=min(step, num_items))
        if parts[p] == num_items:
            part_size = weights[-1] - weights[parts[p - 1]]
            return parts, part_size < bottleneck
        bsum = weights[parts[p] - 1] + bottleneck
    return parts, bsum >= total_weight


def _rb_partition_balanced(weights, num_parts, eps):
    total_weight = weights[-1]
    lower = total_weight / num_parts
    upper = total_weight
    while upper > lower + eps:
        mid = lower + ((upper - lower) / 2)
        parts, success = _lprobe(weights, num_parts, mid)
        if success:
            upper = mid
        else:
            lower = mid + eps
    return upper


def partition_balanced(weights, num_parts, eps=1e-3):
    num_items = len(weights)
    if num_items <= num_parts:
        return partition_uniform(num_items, num_parts)
    weights_ = prefix_sum_inc(weights)
    bottleneck = _rb_partition_balanced(weights_, num_parts, eps=eps)
    parts, success = _lprobe(weights_, num_parts, bottleneck)
    assert success
    return parts


class PartitionedTensor:
    """A tensor split across a group with meta header so it can be re-assembled (activation partitioning / PP)."""

    def __init__(self, tensor, group, partition_meta=None):
        super().__init__()
        self.group = group
        self.num_parts = dist.get_world_size(group=self.group)
        self.rank = dist.get_rank(group=self.group)
        self.orig_size = list(tensor.size())
        self.orig_device = tensor.device
        self.local_data, self.partition = self._partition_tensor(tensor)

    @classmethod
    def from_meta(cls, meta, local_part, group, device=None):
        assert meta.dtype == torch.long
        dummy = torch.ones(dist.get_world_size(group=group))
        part_obj = cls(tensor=dummy, group=group)
        meta = meta.tolist()
        part_obj.orig_size = meta[1:(1 + meta[0])]
        meta = meta[1 + meta[0]:]
        part_obj.orig_device = device or local_part.device
        part_obj.local_data = local_part.detach()
        part_obj.group = group
        assert part_obj.num_parts == meta[0]
        assert part_obj.rank == meta[1]
        part_obj.partition = meta[2:]
        return part_obj

    def _partition_tensor(self, tensor):
        partition = partition_uniform(num_items=tensor.numel(), num_parts=self.num_parts)
        start = partition[self.rank]
        length = partition[self.rank + 1] - start
        tensor_part = tensor.detach().contiguous().view(-1).narrow(0, start=start, length=length).clone()
        return tensor_part, partition

    def full(self, device=None):
        if device is None:
            device = self.orig_device
        full_numel = math.prod(self.full_size())
        flat_tensor = torch.zeros([full_numel], dtype=self.local_data.dtype, device=device)
        partition_tensors = []
        for part_id in range(self.num_parts):
            part_size = self.partition[part_id + 1] - self.partition[part_id]
            buf = flat_tensor.narrow(0, start=self.partition[part_id], length=part_size)
            if part_id == self.rank:
                buf.copy_(self.local_data)
            partition_tensors.append(buf)
        dist.all_gather(partition_tensors, partition_tensors[self.rank], group=self.group)
        return flat_tensor.view(self.full_size()).clone().detach()

    def to_meta(self):
        meta = [len(self.orig_size)] + list(self.orig_size) + [self.num_parts, self.rank] + list(self.partition)
        return torch.LongTensor(data=meta).to(self.orig_device)

    def data(self):
        return self.local_data

    def local_size(self):
        return self.local_data.size()

    def full_size(self):
        return self.orig_size


mem_alloced = 0
mem_cached = 0


def memory_status(msg, print_rank=-1, reset_max=False):
    global mem_alloced, mem_cached
    rank = dist.get_rank()
    if print_rank != -1 and rank != print_rank:
        return
    acc = get_accelerator()
    acc.synchronize()
    if reset_max:
        acc.reset_max_memory_cached()
        acc.reset_max_memory_allocated()
    new_alloced = acc.memory_allocated()
    new_cached = acc.memory_cached()
    delta_alloced = new_alloced - mem_alloced
    delta_cached = new_cached - mem_cached
    mem_cached = new_cached
    mem_alloced = new_alloced
    gb = 1024 ** 3
    print(f"RANK={rank} MEMSTATS", msg, f"device={acc.current_device_name()} current alloc={new_alloced / gb:0.4f}GB "
          f"(delta={delta_alloced / gb:0.4f}GB max={acc.max_memory_allocated() / gb:0.4f}GB) current cache={new_cached / gb:0.4f}GB "
          f"(delta={delta_cached / gb:0.4f}GB max={acc.max_memory_cached() / gb:0.4f}GB)")


def see_memory_usage(message, force=False):
    if not force:
        return
    if dist.is_initialized() and not dist.get_rank() == 0:
        return
    gc.collect()
    acc = get_accelerator()
    logger.info(message)
    logger.info(f"MA {round(acc.memory_allocated() / (1024 ** 3), 2)} GB         Max_MA {round(acc.max_memory_allocated() / (1024 ** 3), 2)} GB "
                f"        CA {round(acc.memory_reserved() / (1024 ** 3), 2)} GB         Max_CA {round(acc.max_memory_reserved() / (1024 ** 3))} GB ")
    try:
        import psutil
        vm_stats = psutil.virtual_memory()
        used_GB = round(((vm_stats.total - vm_stats.available) / (1024 ** 3)), 2)
        logger.info(f"CPU Virtual Memory:  used = {used_GB} GB, percent = {vm_stats.percent}%")
    except ImportError:
        pass
    acc.reset_peak_memory_stats()


def call_to_str(base, *args, **kwargs):
    from .pipe.schedule import call_to_str as c
    return c(base, *args, **kwargs)


def get_only_unique_item(items):
    item_set = set(items)
    if len(item_set) != 1:
        raise RuntimeError(f"expected there to be only one unique element in {items}")
    (unique_item,) = item_set
    return unique_item


def all_gather_dp_groups(groups_flat, partitioned_param_groups, dp_process_group, start_alignment_factor, allgather_bucket_size):
    """All-gather each flat group from its per-rank partitions (parity: runtime/utils.py:956-1002)."""
    for group_id, partitioned_params in enumerate(partitioned_param_groups):
        partition_id = dist.get_rank(group=dp_process_group[group_id])
        dist.all_gather_into_tensor(groups_flat[group_id], partitioned_params[partition_id], group=dp_process_group[group_id])
