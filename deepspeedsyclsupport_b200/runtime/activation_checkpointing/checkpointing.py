"""Activation checkpointing with the reference's options: partition activations across the tensor-parallel group
(all-gather in backward), CPU checkpointing (pinned host copies on a side stream), RNG state tracking for
model-parallel dropout, optional profiling and boundary synchronisation.
Parity: deepspeed/runtime/activation_checkpointing/checkpointing.py:122-246 (RNG tracker), :373-426 (partition),
:484-720 (CheckpointFunction), :724-986 (non-reentrant), :1070-1152 (configure)."""
import contextlib
import copy

import torch

from ... import comm as dist
from ...accelerator import get_accelerator
from ...utils.logging import logger
from ...utils.timer import SynchronizedWallClockTimer as Timers

num_layers = None
mpu = None
mp_rank = None
mp_size = None
mp_group = None
PARTITION_ACTIVATIONS = False
CPU_CHECKPOINT = False
CONTIGUOUS_CHECKPOINTING = False
SYNCHRONIZE = False
PROFILE_TIME = False
deepspeed_checkpointing_enabled = False
timers = None
transport_stream = None
_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


def _dev_rng_get():
    return get_accelerator().get_rng_state()


def _set_cuda_rng_state(new_state, device=-1):
    get_accelerator().set_rng_state(new_state)


class CudaRNGStatesTracker:
    """Named RNG streams so dropout inside tensor-parallel regions differs per rank and elsewhere is identical."""

    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return copy.copy(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception("seed {} already exists".format(seed))
        self.seeds_.add(seed)
        if name in self.states_:
            raise Exception("cuda rng state {} already exists".format(name))
        orig_rng_state = _dev_rng_get()
        get_accelerator().manual_seed(seed)
        self.states_[name] = _dev_rng_get()
        _set_cuda_rng_state(orig_rng_state)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception("cuda rng state {} is not added".format(name))
        orig_cuda_rng_state = _dev_rng_get()
        _set_cuda_rng_state(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = _dev_rng_get()
            _set_cuda_rng_state(orig_cuda_rng_state)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed):
    """default state: same across a tensor-parallel group; model-parallel state: different per tp rank."""
    global mpu
    tp_rank = 0
    if mpu is not None:
        tp_rank = mpu.get_model_parallel_rank() if hasattr(mpu, "get_model_parallel_rank") else 0
    offset = seed + 2718
    model_parallel_seed = offset + tp_rank
    data_parallel_seed = seed
    _CUDA_RNG_STATE_TRACKER.reset()
    get_accelerator().manual_seed(data_parallel_seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, model_parallel_seed)


def model_parallel_reconfigure_tp_seed(seed):
    global mpu
    tp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    model_parallel_seed = seed + 2718 + tp_rank
    with _CUDA_RNG_STATE_TRACKER.fork():
        get_accelerator().manual_seed(model_parallel_seed)


def _mp_world():
    return dist.get_world_size(group=mp_group) if (mp_group is not None and dist.is_initialized()) else 1


def partition_activations(args):
    """Keep only this tp-rank's slice of every floating tensor argument."""
    world, rank = _mp_world(), (dist.get_rank(group=mp_group) if mp_group is not None and dist.is_initialized() else 0)
    out, meta = [], []
    for a in args:
        if torch.is_tensor(a) and a.is_floating_point() and world > 1 and a.numel() % world == 0 and a.numel() > 0:
            n = a.numel() // world
            part = a.detach().contiguous().view(-1).narrow(0, rank * n, n).clone()
            out.append(part)
            meta.append((True, a.shape))
        else:
            out.append(a.detach() if torch.is_tensor(a) else a)
            meta.append((False, None))
    return out, meta


def gather_partitioned_activations(parts, meta):
    world = _mp_world()
    out = []
    for p, (was, shape) in zip(parts, meta):
        if was:
            full = torch.empty(p.numel() * world, dtype=p.dtype, device=p.device)
            dist.all_gather_into_tensor(full, p.contiguous(), group=mp_group)
            out.append(full.view(shape))
        else:
            out.append(p)
    return out


def _to_cpu(t):
    if torch.is_tensor(t) and t.is_cuda:
        buf = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True)
        buf.copy_(t, non_blocking=True)
        return buf
    return t


class CheckpointFunction(torch.autograd.Function):
    """Reentrant checkpoint: forward under no_grad, recompute in backward with restored RNG states."""

    @staticmethod
    def forward(ctx, run_function, all_outputs, *args):
        global timers
        if PROFILE_TIME:
            if timers is None:
                timers = Timers()
            timers("forward").start()
        ctx.run_function = run_function
        ctx.fwd_cpu_rng_state = torch.get_rng_state()
        ctx.fwd_dev_rng_state = _dev_rng_get() if get_accelerator().device_name() == "cuda" else None
        ctx.fwd_tracker_states = get_cuda_rng_tracker().get_states()
        ctx.tensor_idx = [i for i, a in enumerate(args) if torch.is_tensor(a)]
        ctx.non_tensor = {i: a for i, a in enumerate(args) if not torch.is_tensor(a)}
        tensors = [args[i] for i in ctx.tensor_idx]
        ctx.requires = [t.requires_grad for t in tensors]
        if PARTITION_ACTIVATIONS:
            saved, ctx.part_meta = partition_activations(tensors)
        else:
            saved, ctx.part_meta = [t.detach() for t in tensors], None
        ctx.on_cpu = False
        if CPU_CHECKPOINT:
            saved = [_to_cpu(t) for t in saved]
            ctx.on_cpu = True
            ctx.devices = [t.device for t in tensors]
        with torch.no_grad():
            outputs = run_function(*args)
        ctx.save_for_backward(*saved)
        if SYNCHRONIZE:
            get_accelerator().synchronize()
        if PROFILE_TIME:
            timers("forward").stop()
            timers.log(["forward"])
        if torch.is_tensor(outputs):
            non_grad = []
            all_outputs += [outputs]
            return outputs
        all_outputs += list(outputs)
        return tuple(outputs)

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")
        saved = list(ctx.saved_tensors)
        if ctx.on_cpu:
            saved = [t.to(d, non_blocking=True) if torch.is_tensor(t) else t for t, d in zip(saved, ctx.devices)]
        if ctx.part_meta is not None:
            saved = gather_partitioned_activations(saved, ctx.part_meta)
        n_args = len(ctx.tensor_idx) + len(ctx.non_tensor)
        args = [None] * n_args
        detached = []
        for i, t, rg in zip(ctx.tensor_idx, saved, ctx.requires):
            t = t.detach()
            t.requires_grad = rg
            args[i] = t
            detached.append(t)
        for i, a in ctx.non_tensor.items():
            args[i] = a
        bwd_cpu_rng_state = torch.get_rng_state()
        bwd_dev_rng_state = _dev_rng_get() if ctx.fwd_dev_rng_state is not None else None
        bwd_tracker = get_cuda_rng_tracker().get_states()
        torch.set_rng_state(ctx.fwd_cpu_rng_state)
        if ctx.fwd_dev_rng_state is not None:
            _set_cuda_rng_state(ctx.fwd_dev_rng_state)
        get_cuda_rng_tracker().set_states(ctx.fwd_tracker_states)
        with torch.enable_grad():
            outputs = ctx.run_function(*args)
        torch.set_rng_state(bwd_cpu_rng_state)
        if bwd_dev_rng_state is not None:
            _set_cuda_rng_state(bwd_dev_rng_state)
        get_cuda_rng_tracker().set_states(bwd_tracker)
        if isinstance(outputs, torch.Tensor):
            outputs = (outputs, )
        output_tensors, grad_tensors = [], []
        for out, grad in zip(outputs, grads):
            if torch.is_tensor(out) and out.requires_grad:
                output_tensors.append(out)
                grad_tensors.append(grad)
        if output_tensors:
            torch.autograd.backward(output_tensors, grad_tensors)
        ret = [None, None]
        k = 0
        for i in range(n_args):
            if i in ctx.non_tensor:
                ret.append(None)
            else:
                ret.append(detached[k].grad if detached[k].requires_grad else None)
                k += 1
        return tuple(ret)


def non_reentrant_checkpoint(function, *args):
    """Saved-tensor-hook based variant (no autograd.Function re-entry)."""
    from torch.utils.checkpoint import checkpoint as torch_checkpoint
    return torch_checkpoint(function, *args, use_reentrant=False)


def checkpoint(function, *args):
    """Checkpoint a model or part of the model (drop-in for torch.utils.checkpoint.checkpoint)."""
    all_outputs = []
    CheckpointFunction.apply(function, all_outputs, *args)
    if len(all_outputs) == 1:
        return all_outputs[0]
    return tuple(all_outputs)


def partition_activations_in_checkpoint(partition_activation):
    global PARTITION_ACTIVATIONS
    PARTITION_ACTIVATIONS = partition_activation
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"**************Partition Activations {PARTITION_ACTIVATIONS}************")


def set_num_layers(nlayers):
    global num_layers
    num_layers = nlayers


def reset():
    """Release contiguous buffers between iterations (no-op: this implementation allocates per checkpoint)."""
    return


def _configure_using_config_file(config, mpu=None):
    global num_layers, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME
    from ..config import DeepSpeedConfig
    config = DeepSpeedConfig(config, mpu=mpu).activation_checkpointing_config
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(config.model_dump())
    PARTITION_ACTIVATIONS = config.partition_activations
    CONTIGUOUS_CHECKPOINTING = config.contiguous_memory_optimization
    num_layers = config.number_checkpoints
    CPU_CHECKPOINT = config.cpu_checkpointing
    SYNCHRONIZE = config.synchronize_checkpoint_boundary
    PROFILE_TIME = config.profile


def _configure_defaults():
    global mpu, num_layers, deepspeed_checkpointing_enabled, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING
    global CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME
    PARTITION_ACTIVATIONS = False
    CONTIGUOUS_CHECKPOINTING = False
    num_layers = False
    CPU_CHECKPOINT = False
    SYNCHRONIZE = False
    PROFILE_TIME = False
    deepspeed_checkpointing_enabled = True


def configure(mpu_, deepspeed_config=None, partition_activations=None, contiguous_checkpointing=None, num_checkpoints=None,
              checkpoint_in_cpu=None, synchronize=None, profile=None):
    global mpu, num_layers, deepspeed_checkpointing_enabled, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING
    global CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME, mp_group, mp_rank, mp_size
    _configure_defaults()
    if mpu_ is not None:
        mpu = mpu_
        mp_group = mpu.get_model_parallel_group() if hasattr(mpu, "get_model_parallel_group") else None
    if deepspeed_config is not None:
        _configure_using_config_file(deepspeed_config, mpu=mpu)
    if partition_activations is not None:
        PARTITION_ACTIVATIONS = partition_activations
    if contiguous_checkpointing is not None:
        CONTIGUOUS_CHECKPOINTING = contiguous_checkpointing
    if num_checkpoints is not None:
        num_layers = num_checkpoints
    if checkpoint_in_cpu is not None:
        CPU_CHECKPOINT = checkpoint_in_cpu
    if synchronize is not None:
        SYNCHRONIZE = synchronize
    if profile is not None:
        PROFILE_TIME = profile
    if CONTIGUOUS_CHECKPOINTING:
        assert PARTITION_ACTIVATIONS, "Contiguous Checkpointing is only available with partitioned activations. Set partitioned activations to true in deepspeed config"
    if CONTIGUOUS_CHECKPOINTING:
        assert num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"


def is_configured():
    return deepspeed_checkpointing_enabled
