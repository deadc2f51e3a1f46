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


# ---------------------------------------------------------------------------------------------------------------------
# Overflow check, tensor-list norms / clipping, CUDA-graph helper and small shims.
# ---------------------------------------------------------------------------------------------------------------------
_graphs = {}


def graph_process(replay_first_step, func, *args, **kwargs):
    """Capture ``func(*args, **kwargs)`` in a CUDA graph on first use and replay it afterwards
    (parity: runtime/utils.py:53-68).  The arguments must be static tensors; on a host-only build the
    function simply runs."""
    if not torch.cuda.is_available():
        func(*args, **kwargs)
        return
    key = func.__name__ + "_graph"
    first = key not in _graphs
    if first:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up outside capture so lazy inits do not land in the graph
            func(*args, **kwargs)
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            func(*args, **kwargs)
        _graphs[key] = g
    if not first or replay_first_step:
        _graphs[key].replay()


def copy_to_device(item, device, criterion_func):
    """Like ``move_to_device`` but always makes a fresh copy of matching tensors
    (parity: runtime/utils.py:158-181)."""
    if criterion_func(item):
        return item.to(device, copy=True)
    if isinstance(item, list):
        return [copy_to_device(v, device, criterion_func) for v in item]
    if isinstance(item, tuple):
        return tuple(copy_to_device(v, device, criterion_func) for v in item)
    if isinstance(item, dict):
        return {k: copy_to_device(v, device, criterion_func) for k, v in item.items()}
    return item


class CheckOverflow:
    """inf/nan detection over gradients with the result agreed on by every rank that shares them
    (parity: runtime/utils.py:208-322).  One fused ``_foreach_norm`` pass replaces the per-tensor sum."""

    def __init__(self, param_groups=None, mpu=None, zero_reduce_scatter=False, deepspeed=None):
        self.mpu = mpu
        self.params = [] if param_groups else None
        self.zero_reduce_scatter = zero_reduce_scatter
        self.deepspeed = deepspeed
        self.has_moe_params = False
        if param_groups:
            for group in param_groups:
                for p in group:
                    self.params.append(p)
                    if getattr(p, "allreduce", True) is False:
                        self.has_moe_params = True

    def _agree(self, overflow, has_moe_params=None):
        flag = torch.tensor([float(overflow)], device=get_accelerator().current_device_name()
                            if torch.cuda.is_available() else "cpu")
        if not dist.is_initialized():
            return bool(flag.item())
        if has_moe_params is None:
            has_moe_params = self.has_moe_params
        if has_moe_params:
            from ..utils import groups
            try:
                dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=groups._get_max_expert_parallel_group())
            except Exception:
                pass
        if self.zero_reduce_scatter:
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=dist.get_world_group())
        elif self.mpu is not None:
            if self.deepspeed is not None and getattr(self.deepspeed, "pipeline_enable_backward_allreduce", True) is False:
                dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.mpu.get_data_parallel_group())
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.mpu.get_model_parallel_group())
        elif self.deepspeed is not None and getattr(self.deepspeed, "pipeline_enable_backward_allreduce", True) is False:
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=dist.get_world_group())
        return bool(flag.item())

    def check_using_norm(self, norm_group, reduce_overflow=True):
        overflow = -1 in norm_group
        if not reduce_overflow:
            return bool(overflow)
        return self._agree(overflow, has_moe_params=False)

    def check(self, param_groups=None):
        if param_groups is None:
            assert self.params is not None, "CheckOverflow.check: no parameters given at construction or call"
            params = self.params
            has_moe = self.has_moe_params
        else:
            assert param_groups, "CheckOverflow.check: empty param_groups"
            params = [p for g in param_groups for p in g]
            has_moe = any(getattr(p, "allreduce", True) is False for p in params)
        return self.has_overflow(params, has_moe_params=has_moe)

    def has_overflow_serial(self, params):
        grads = [p.grad.data for p in params if p.grad is not None]
        if not grads:
            return False
        dense = [g.float() if g.dtype not in (torch.float32, torch.float16, torch.bfloat16) else g for g in grads
                 if not g.is_sparse]
        sparse = [g.coalesce().values() for g in grads if g.is_sparse]
        norms = torch._foreach_norm(dense + sparse) if (dense or sparse) else []
        total = torch.stack([n.float() for n in norms]).sum() if norms else torch.zeros(())
        return not bool(torch.isfinite(total).item())

    def has_overflow(self, params, has_moe_params=None):
        return self._agree(self.has_overflow_serial(params), has_moe_params=has_moe_params)

    @staticmethod
    def _has_inf_or_nan(x, i=None):
        return not bool(torch.isfinite(x.float().sum()).item())


def empty_cache():
    """Release cached device blocks and reset the peak counters (parity: runtime/utils.py:776-778)."""
    get_accelerator().empty_cache()
    if torch.cuda.is_available():
        get_accelerator().reset_peak_memory_stats()


def get_global_norm_of_tensors(input_tensors, norm_type=2, mpu=None, use_graph=False):
    """p-norm of a list of tensors viewed as one vector, reduced over the model-parallel group; -1 when the
    result is inf/nan (parity: runtime/utils.py:855-905).  A single multi-tensor norm launch covers the list, so
    ``use_graph`` needs no separate capture and is accepted for signature parity."""
    input_tensors = list(input_tensors)
    assert all(torch.is_tensor(t) for t in input_tensors), "expected an iterable of tensors"
    norm_type = float(norm_type)
    if not input_tensors:
        return 0.0
    device = input_tensors[0].device
    if norm_type == math.inf:
        total = torch.stack([t.detach().abs().max().float() for t in input_tensors]).max().reshape(1).to(device)
        if mpu is not None:
            dist.all_reduce(total, op=dist.ReduceOp.MAX, group=mpu.get_model_parallel_group())
        total_norm = total.item()
    else:
        per = torch._foreach_norm([t.detach().float() for t in input_tensors], norm_type)
        total = (torch.stack(per) ** norm_type).sum().reshape(1)
        if mpu is not None:
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=mpu.get_model_parallel_group())
        total_norm = total.item() ** (1.0 / norm_type)
    if math.isinf(total_norm) or math.isnan(total_norm):
        total_norm = -1
    return total_norm


def clip_tensors_by_global_norm(input_tensors, max_norm=1.0, global_norm=None, mpu=None, eps=1e-6, use_graph=False):
    """Scale the tensors in place so their global norm is at most ``max_norm``; returns the norm before clipping
    (parity: runtime/utils.py:908-939)."""
    input_tensors = list(input_tensors)
    if global_norm is None:
        global_norm = get_global_norm_of_tensors(input_tensors, mpu=mpu, use_graph=use_graph)
    clip_coef = max_norm / (global_norm + eps)
    if clip_coef < 1 and input_tensors:
        torch._foreach_mul_([t.detach() for t in input_tensors], clip_coef)
    return global_norm


def align_dense_tensors(tensor_list, alignment):
    """Append a zero pad tensor so the total element count is a multiple of ``alignment``
    (parity: runtime/utils.py:942-953)."""
    numel = sum(t.numel() for t in tensor_list)
    rem = numel % alignment
    if rem == 0:
        return tensor_list
    pad = torch.zeros(alignment - rem, device=tensor_list[0].device, dtype=tensor_list[0].dtype)
    return list(tensor_list) + [pad]


def all_gather_into_tensor_dp_groups(groups_flat, partitioned_param_groups, dp_process_group):
    """One ``all_gather_into_tensor`` per flat group, each rank contributing its own partition
    (parity: runtime/utils.py:956-961)."""
    for group_id, (flat, partitions) in enumerate(zip(groups_flat, partitioned_param_groups)):
        rank = dist.get_rank(group=dp_process_group[group_id])
        if dist.get_world_size(group=dp_process_group[group_id]) <= 1:
            continue
        dist.all_gather_into_tensor(flat, partitions[rank], group=dp_process_group[group_id])


class TLinear(torch.nn.Linear):
    """Linear built from a layer whose weight is held as ``[in, out]`` (the layout hybrid-engine containers keep):
    the weight is transposed once at construction, the forward is a plain linear (parity: runtime/utils.py:1005-1021)."""

    def __init__(self, orig_layer, name=""):
        self.name = name
        w = orig_layer.weight
        torch.nn.Module.__init__(self)
        self.in_features, self.out_features = w.shape[0], w.shape[1]
        self.weight = torch.nn.Parameter(w.data.t().contiguous(), requires_grad=w.requires_grad)
        self.bias = orig_layer.bias

    def forward(self, input):
        return torch.nn.functional.linear(input, self.weight, self.bias)


def get_inactive_params(param_list):
    """ZeRO-3 parameters of ``param_list`` whose full tensor is not resident (parity: runtime/utils.py:1024-1027)."""
    from .zero.partition_parameters import ZeroParamStatus
    return [p for p in param_list if hasattr(p, "ds_id") and p.ds_status == ZeroParamStatus.NOT_AVAILABLE]


def required_torch_version(min_version=None, max_version=None):
    """True when the installed torch ``major.minor`` lies in ``[min_version, max_version]``
    (parity: runtime/utils.py:1030-1041)."""
    assert min_version or max_version, "give min_version, max_version or both"
    from packaging import version as pkg_version
    parts = torch.__version__.split("+")[0].split(".")
    cur = pkg_version.parse(f"{parts[0]}.{parts[1]}")
    if min_version and cur < pkg_version.parse(str(min_version)):
        return False
    if max_version and cur > pkg_version.parse(str(max_version)):
        return False
    return True
