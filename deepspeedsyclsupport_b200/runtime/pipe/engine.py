"""PipelineEngine: runs a TrainSchedule / InferenceSchedule instruction stream on one pipeline stage.

How it is built (behavioural parity with deepspeed/runtime/pipe/engine.py:55-1372 -- `train_batch`, `eval_batch`, the
instruction set of runtime/pipe/schedule.py, tied-weight and data-parallel gradient reduction, loss aggregation -- but not
its mechanics):

* **Slots, not buffer lists.**  Every in-flight micro-batch owns a `_Slot` (inputs, labels, outputs, incoming output
  gradients) together with the communication it is waiting for.
* **Asynchronous stage-to-stage transfers.**  Receives are posted (`irecv`) when the schedule says so, straight into freshly
  allocated tensors, and are only waited for by the instruction that consumes them (`ForwardPass` / `BackwardPass`).  Sends
  are `isend`s whose handles stay with the slot; the tensors they read are kept alive by the slot until the handle has
  completed (checked when the slot is recycled and when the schedule drains).  On NCCL `wait()` orders the compute stream
  after the transfer without blocking the host, so the copy of micro-batch n+1 runs under the math of micro-batch n.
* **One descriptor message per wire.**  The first transfer on a wire (activations to the next stage, gradients need none:
  they mirror the outputs) is preceded by a single int64 descriptor: [kind, count, then per tensor dtype id, rank, dims].
  The receiver caches it; `reset_activation_shape()` forgets it on both sides.
* Pipeline-stage transfers stay on NCCL send/recv (BASELINE north star); ZeRO stages 0/1 compose with pipeline parallelism.
"""
import os

import torch

from ... import comm as dist
from ...accelerator import get_accelerator
from ...utils.logging import logger
from ..dataloader import RepeatingLoader
from ..engine import DeepSpeedEngine
from . import p2p, schedule
from .module import PipelineError, PipelineModule

TARGET_ID = -2
LOG_STAGE = -2
DATA_PARALLEL_ID = -2

_KIND_TENSOR, _KIND_LIST, _KIND_TUPLE = 0, 1, 2
_DESC_LEN = 256  # int64 words; enough for dozens of tensors of ordinary rank


def is_even(number):
    return number % 2 == 0


def _as_tuple(x):
    return (x, ) if torch.is_tensor(x) else tuple(x)


class _Slot:
    """State of one in-flight micro-batch on this stage."""
    __slots__ = ("inputs", "labels", "outputs", "out_grads", "pending_in", "pending_grad", "sends")

    def __init__(self):
        self.inputs = self.labels = self.outputs = self.out_grads = None
        self.pending_in, self.pending_grad, self.sends = [], [], []

    def settle_sends(self):
        for work, _keepalive in self.sends:
            if work is not None:
                work.wait()
        self.sends = []


class PipelineEngine(DeepSpeedEngine):
    ID_TO_DTYPE = [torch.float32, torch.float64, torch.complex64, torch.complex128, torch.float16, torch.bfloat16, torch.uint8,
                   torch.int8, torch.int16, torch.int32, torch.int64, torch.bool]
    DTYPE_TO_ID = {dtype: id_ for id_, dtype in enumerate(ID_TO_DTYPE)}

    def __init__(self, has_bool_tensors=False, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        grid = self.grid = self.module._grid
        # geometry
        self.global_rank = grid.get_global_rank()
        self.num_stages, self.stage_id = grid.pipe_parallel_size, grid.get_stage_id()
        self.prev_stage, self.next_stage = self.stage_id - 1, self.stage_id + 1
        self._first_stage, self._last_stage = self.stage_id == 0, self.stage_id == self.num_stages - 1
        self.is_pipe_parallel = self.num_stages > 1
        self.is_data_parallel = grid.data_parallel_size > 1
        self.is_model_parallel = grid.model_parallel_size > 1
        pipe_cfg = self._config.pipeline
        self.is_pipe_partitioned = self.is_model_parallel and pipe_cfg["pipe_partitioned"]
        self.is_grad_partitioned = self.is_model_parallel and pipe_cfg["grad_partitioned"]
        # batch shape
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        assert self.dp_world_size == grid.data_parallel_size
        assert self.train_batch_size() == self.micro_batch_size * self.micro_batches * grid.data_parallel_size
        # behaviour switches kept under the reference's names
        self.enable_backward_allreduce = False  # the engine's own per-step all-reduce is replaced by ReduceGrads
        self.pipeline_enable_backward_allreduce = True
        self.has_bool_tensors = has_bool_tensors
        self.eval_return_logits = False
        self.using_bf16_optimizer = False
        self.log_batch_step_id = -1
        self.batch_timer = None
        self.has_attention_mask = type(self.module).__name__ == "GPT2ModelPipe"
        self._force_grad_boundary = False
        self._compute_loss = True
        # data
        self.data_iterator = self.batch_fn = None
        if self.training_data:
            self._build_data_iter(self.training_data)
        # per-schedule state
        self._slots = []
        self._in_desc = None           # cached descriptor of what the previous stage sends
        self._out_desc_sent = False
        self.outputs = None
        self.fwd_outputs = []
        self.loss = torch.zeros((), device=self.device)
        self.total_loss = None
        self.agg_loss = torch.zeros((), device=self.device)
        self.dp_group_loss = torch.zeros((), device=self.device)
        if pipe_cfg["activation_checkpoint_interval"] > 0:
            self.module.activation_checkpoint_interval = pipe_cfg["activation_checkpoint_interval"]
        self.module.checkpoint_parallel_write_pipeline = self._config.checkpoint_parallel_write_pipeline
        if not hasattr(type(self.module), "loss_fn"):
            type(self.module).loss_fn = self.module.loss_fn
        if self.is_pipe_parallel:
            p2p.init_process_groups(grid)

    # ------------------------------------------------------------------ small public surface
    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def is_first_stage(self):
        """True on the stage that reads the inputs (reference: pipe/engine.py:509)."""
        return self._first_stage

    def is_last_stage(self):
        return self._last_stage

    is_first_stage_fn, is_last_stage_fn = is_first_stage, is_last_stage

    def is_gradient_accumulation_boundary(self):
        return self._force_grad_boundary

    def set_train_batch_size(self, train_batch_size):
        super().set_train_batch_size(train_batch_size)
        self.micro_batches = self.gradient_accumulation_steps()

    def reset_activation_shape(self):
        """Forget the cached wire descriptors (call on every stage when the activation shape changes)."""
        self._in_desc, self._out_desc_sent = None, False

    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    backward = step = forward

    # ------------------------------------------------------------------ data
    def _build_data_iter(self, dataset):
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        self.set_dataloader(RepeatingLoader(self.deepspeed_io(dataset, data_sampler=sampler)))

    def _reads_data(self):
        return self._first_stage or self._last_stage  # inputs on the first stage, labels on the last

    def set_dataloader(self, loader):
        if self._reads_data():
            self.training_dataloader, self.data_iterator = loader, iter(loader)

    def set_dataiterator(self, iterator):
        if self._reads_data():
            self.training_dataloader, self.data_iterator = None, iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def _next_batch(self):
        batch = next(self.data_iterator) if self.data_iterator is not None else None
        return self.batch_fn(batch) if self.batch_fn else batch

    # ------------------------------------------------------------------ entry points
    def train_batch(self, data_iter=None):
        """One global batch: `micro_batches` forward/backward passes through the pipeline, gradient reduction, optimizer
        step.  Returns the batch-mean loss, valid on every rank."""
        if not torch.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss, self._compute_loss = None, True
        self._run(schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id))
        self.agg_train_loss = self._aggregate_total_loss()
        if self.global_rank == 0:
            if self.global_steps % self.steps_per_print() == 0:
                print(f"steps: {self.global_steps} loss: {self.agg_train_loss:0.4f} ")
            if self.monitor.enabled:
                self.monitor.write_events([("Train/Samples/train_loss", self.agg_train_loss.mean().item(), self.global_samples)])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, compute_loss=True, reduce_output="avg", bcast_loss=True, num_micro_batches=None):
        self.eval_return_logits, self._compute_loss = return_logits, compute_loss
        self.module.eval()
        saved_iterator = self.data_iterator
        self.set_dataiterator(data_iter)
        n = self.micro_batches if num_micro_batches is None else num_micro_batches
        dist.barrier()
        with torch.no_grad():
            self._run(schedule.InferenceSchedule(micro_batches=n, stages=self.num_stages, stage_id=self.stage_id))
        result = self._reduce_outputs(self.fwd_outputs, reduce=reduce_output, micro_batches=n) if self._last_stage else None
        if compute_loss and (bcast_loss or self.monitor.enabled):
            result = self._bcast_pipe_scalar(result)
        self.set_dataiterator(saved_iterator)
        if not return_logits:
            return result
        logits, self.outputs = self.outputs, None
        return result, logits

    # ------------------------------------------------------------------ schedule interpreter
    def _run(self, sched):
        while len(self._slots) < sched.num_pipe_buffers():
            self._slots.append(_Slot())
        self.fwd_outputs = []
        handlers = self._handlers()
        for step_cmds in sched:
            for cmd in step_cmds:
                handler = handlers.get(type(cmd))
                if handler is None:
                    raise RuntimeError(f"{type(self).__name__} does not understand instruction {cmd!r}")
                handler(**cmd.kwargs)
        for slot in self._slots:  # drain: nothing of this batch may still be in flight when the next one starts
            slot.settle_sends()
            slot.inputs = slot.outputs = slot.out_grads = None

    def _handlers(self):
        return {
            schedule.LoadMicroBatch: self._exec_load_micro_batch,
            schedule.RecvActivation: self._exec_recv_activations,
            schedule.ForwardPass: self._exec_forward_pass,
            schedule.SendActivation: self._exec_send_activations,
            schedule.RecvGrad: self._exec_recv_grads,
            schedule.BackwardPass: self._exec_backward_pass,
            schedule.SendGrad: self._exec_send_grads,
            schedule.ReduceTiedGrads: self._exec_reduce_tied_grads,
            schedule.ReduceGrads: self._exec_reduce_grads,
            schedule.OptimizerStep: self._exec_optimizer_step,
        }

    # ---- wire descriptors
    def _describe(self, payload):
        kind = _KIND_TENSOR if torch.is_tensor(payload) else (_KIND_TUPLE if isinstance(payload, tuple) else _KIND_LIST)
        words = [kind, len(_as_tuple(payload))]
        for t in _as_tuple(payload):
            assert torch.is_tensor(t), f"pipeline stages exchange tensors, got {type(t)}"
            words += [self.DTYPE_TO_ID[t.dtype], t.dim(), *t.shape]
        assert len(words) <= _DESC_LEN, "activation descriptor too long"
        return torch.tensor(words + [0] * (_DESC_LEN - len(words)), dtype=torch.long, device=self.device)

    def _parse(self, desc):
        words = desc.tolist()
        kind, count, pos, specs = words[0], words[1], 2, []
        for _ in range(count):
            dtype, rank = self.ID_TO_DTYPE[words[pos]], words[pos + 1]
            specs.append((tuple(words[pos + 2:pos + 2 + rank]), dtype))
            pos += 2 + rank
        return kind, specs

    def _blank(self, specs):
        return [torch.empty(shape, dtype=dtype, device=self.device) for shape, dtype in specs]

    # ---- instructions: data in
    def _exec_load_micro_batch(self, buffer_id):
        slot = self._slots[buffer_id]
        slot.settle_sends()
        batch = self._next_batch()

        def to_device(x, want_grad):
            y = x.clone().detach().to(self.device)
            y.requires_grad = want_grad and y.is_floating_point()
            return y

        if self._first_stage:
            x = batch[0]
            assert torch.is_tensor(x) or isinstance(x, (tuple, list)), f"unsupported input type {type(x)}"
            slot.inputs = to_device(x, True) if torch.is_tensor(x) else tuple(to_device(t, True) for t in x)
        if self._last_stage:
            y = batch[1]
            if torch.is_tensor(y):
                y = y.to(self.device)
            elif isinstance(y, (tuple, list)):
                y = tuple(t.to(self.device).detach() for t in y)
            slot.labels = y

    def _exec_recv_activations(self, buffer_id):
        slot = self._slots[buffer_id]
        slot.settle_sends()
        if self._in_desc is None:
            desc = torch.zeros(_DESC_LEN, dtype=torch.long, device=self.device)
            p2p.recv(desc, self.prev_stage)
            self._in_desc = self._parse(desc)
        kind, specs = self._in_desc
        tensors = self._blank(specs)
        slot.pending_in = [p2p.recv(t, self.prev_stage, async_op=True) for t in tensors]
        slot.inputs = tensors[0] if kind == _KIND_TENSOR else tuple(tensors)

    # ---- instructions: compute
    def _exec_forward_pass(self, buffer_id):
        self.tput_timer.start()
        slot = self._slots[buffer_id]
        for work in slot.pending_in:
            work.wait()
        slot.pending_in = []
        inputs = slot.inputs
        if not self._first_stage:  # received tensors become graph leaves whose .grad goes back upstream
            for t in _as_tuple(inputs):
                t.requires_grad = t.is_floating_point() and torch.is_grad_enabled()
        for t in _as_tuple(inputs):
            t.grad = None
        single = torch.is_tensor(inputs) or len(inputs) == 1
        slot.outputs = outputs = self.module.forward(inputs if torch.is_tensor(inputs) else (inputs[0] if single else inputs))
        if not self._last_stage:
            return
        if self._compute_loss and self.module.loss_fn is not None:
            self.loss = self.module.loss_fn(outputs, slot.labels)
        else:
            self.loss = outputs
        if self.eval_return_logits:
            self.outputs = outputs
        parts = [self.loss.detach()] if torch.is_tensor(self.loss) else [l.detach() for l in self.loss]
        self.fwd_outputs.append(parts[0] if torch.is_tensor(self.loss) else parts)
        if self.total_loss is None:
            self.total_loss = torch.zeros_like(parts[0]) if torch.is_tensor(self.loss) else [torch.zeros_like(p) for p in parts]
        if torch.is_tensor(self.loss):
            self.total_loss += parts[0]
        else:
            for acc, p in zip(self.total_loss, parts):
                acc += p

    def _exec_backward_pass(self, buffer_id):
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        slot = self._slots[buffer_id]
        if self._last_stage:
            loss = self._scale_loss_by_gas(self.loss.float()) if self.gradient_accumulation_steps() > 1 else self.loss
            self.optimizer.backward(loss)
            self.optimizer.micro_step_id += 1
            return
        for work in slot.pending_grad:
            work.wait()
        slot.pending_grad = []
        slot.settle_sends()  # the activations this stage sent downstream have long been consumed
        roots = [t for t in _as_tuple(slot.outputs) if t.is_floating_point()]
        assert len(roots) == len(slot.out_grads), "one incoming gradient per floating-point output"
        torch.autograd.backward(tensors=roots, grad_tensors=slot.out_grads)
        slot.outputs = slot.out_grads = None

    # ---- instructions: data out
    def _exec_send_activations(self, buffer_id):
        slot = self._slots[buffer_id]
        payload = slot.outputs
        if not (torch.is_tensor(payload) or isinstance(payload, tuple)):
            raise NotImplementedError(f"Could not send output of type {type(payload)}")
        if not self._out_desc_sent:
            p2p.send(self._describe(payload), self.next_stage)
            self._out_desc_sent = True
        for t in _as_tuple(payload):
            slot.sends.append((p2p.send(t, self.next_stage, async_op=True), t))

    def _exec_recv_grads(self, buffer_id):
        slot = self._slots[buffer_id]
        specs = [(tuple(t.shape), t.dtype) for t in _as_tuple(slot.outputs) if t.is_floating_point()]
        slot.out_grads = self._blank(specs)
        slot.pending_grad = [p2p.recv(g, self.next_stage, async_op=True) for g in slot.out_grads]

    def _exec_send_grads(self, buffer_id):
        slot = self._slots[buffer_id]
        for t in _as_tuple(slot.inputs):
            if not t.is_floating_point():
                assert t.grad is None
                continue
            assert t.grad is not None, "a floating-point pipeline input received no gradient"
            slot.sends.append((p2p.send(t.grad, self.prev_stage, async_op=True), t.grad))
        slot.inputs = None

    # ---- instructions: end of batch
    def _exec_reduce_tied_grads(self):
        """Gradients of weights shared between stages are summed over the stages that hold a copy."""
        for weight, group in self.module.get_tied_weights_and_groups():
            zero_opt = getattr(weight, "ds_zero_optimizer", None)
            grad = zero_opt.main_grad_view(weight) if zero_opt is not None else weight.grad
            if grad is not None:
                dist.all_reduce(grad, group=group)

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        try:
            if self.pipeline_enable_backward_allreduce:
                self.optimizer.end_micro_step(True)
        finally:
            self._force_grad_boundary = False

    def _exec_optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        try:
            self._take_model_step(lr_kwargs)
        finally:
            self._force_grad_boundary = False
        self.micro_steps += self.micro_batches

    # ------------------------------------------------------------------ losses / outputs
    def _last_stage_rank(self):
        rank = self.grid.stage_to_global(self.num_stages - 1)
        assert rank in self.grid.pp_group
        return rank

    def _reduce_outputs(self, outputs, reduce="avg", reduce_dp=True, micro_batches=None):
        if reduce is None:
            return outputs
        if reduce.lower() != "avg":
            raise NotImplementedError(f"reduction type {reduce} not supported.")
        if torch.is_tensor(outputs[0]):
            mean = sum(outputs) / (micro_batches or self.micro_batches)
            if reduce_dp and self.is_data_parallel:
                dist.all_reduce(mean, group=self.mpu.get_data_parallel_group())
                mean /= self.dp_world_size
            return mean
        assert isinstance(outputs, (list, tuple))
        sums = [torch.zeros_like(o) for o in outputs[0]]
        for per_micro_batch in outputs:
            for acc, o in zip(sums, per_micro_batch):
                acc += o
        return sums

    def _bcast_pipe_scalar(self, data, src_rank=None, dtype=torch.float32):
        src = self._last_stage_rank() if src_rank is None else src_rank
        if self.global_rank == src:
            value = data.clone().detach().to(device=self.device, dtype=dtype)
        else:
            value = torch.zeros(1, dtype=dtype, device=self.device)
        dist.broadcast(tensor=value, src=src, group=self.mpu.get_pipe_parallel_group())
        return value

    def _aggregate_total_loss(self):
        """[this replica's mean loss, mean over data-parallel replicas], computed on the last stage, broadcast up the pipe."""
        pair = torch.zeros(2, dtype=torch.float32, device=self.device)
        if self._last_stage:
            mine = self._scale_loss_by_gas(self.total_loss).clone().detach().float()
            everyone = mine.clone()
            if self.is_data_parallel:
                dist.all_reduce(everyone, group=self.mpu.get_data_parallel_group())
                everyone /= self.dp_world_size
            pair = torch.stack([mine, everyone])
        if self.is_pipe_parallel:
            dist.broadcast(tensor=pair, src=self._last_stage_rank(), group=self.mpu.get_pipe_parallel_group())
        self.dp_group_loss = pair[0].clone().detach()
        return pair[1].clone().detach()

    # ------------------------------------------------------------------ logging helpers
    def log_for_device(self, *msg):
        """Print on the stage / data-parallel replica selected by DSB200_PIPE_LOG_STAGE / DSB200_PIPE_LOG_DP (-1 = all)."""
        want_stage = int(os.environ.get("DSB200_PIPE_LOG_STAGE", str(LOG_STAGE)))
        want_dp = int(os.environ.get("DSB200_PIPE_LOG_DP", "-1"))
        if want_stage in (-1, self.stage_id) and want_dp in (-1, self.grid.data_parallel_id):
            print(f"RANK={dist.get_rank()} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} STEP-ID={self.global_steps}", *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.global_steps % self.steps_per_print() == 0:
            print(*msg)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        """Allocated / reserved device memory and the change since the previous call, on one rank (or all with -1)."""
        acc = get_accelerator()
        if not acc.is_available() or print_rank not in (-1, self.global_rank):
            return
        acc.synchronize()
        if reset_max:
            acc.reset_max_memory_allocated()
        now = (acc.memory_allocated(), acc.memory_reserved() if hasattr(acc, "memory_reserved") else 0)
        before, self._mem_status_prev = getattr(self, "_mem_status_prev", (0, 0)), now
        gib = 2.0 ** 30
        print(f"RANK={self.global_rank} STAGE={self.stage_id} STEP={self.global_steps} MEMSTATS {msg} "
              f"alloc={now[0] / gib:.4f}GB (delta {(now[0] - before[0]) / gib:+.4f}GB, max {acc.max_memory_allocated() / gib:.4f}GB) "
              f"reserved={now[1] / gib:.4f}GB (delta {(now[1] - before[1]) / gib:+.4f}GB)", flush=True)

    # ------------------------------------------------------------------ checkpoints: one file per layer, written by the module
    _curr_ckpt_path = None

    def module_state_dict(self, exclude_frozen_parameters=False):
        """PipelineModule.save_state_dict writes layer_XX-model_YY-model_states.pt files; nothing goes into the engine's dict."""
        assert self._curr_ckpt_path is not None, "PipelineEngine expects module_state_dict() to be called from save_checkpoint()"
        self.module.save_state_dict(self._curr_ckpt_path, checkpoint_engine=self.checkpoint_engine,
                                    exclude_frozen_params=exclude_frozen_parameters)
        return None

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False, state_dict=None):
        assert custom_load_fn is None, "custom_load_fn not supported w. pipeline parallelism"
        sd = checkpoint["module"] if checkpoint is not None else state_dict
        if sd is not None and not isinstance(sd, str):
            return super().load_module_state_dict(state_dict=sd, strict=strict)
        self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict, checkpoint_engine=self.checkpoint_engine)
        if hasattr(self.optimizer, "refresh_from_params"):
            self.optimizer.refresh_from_params()

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True, exclude_frozen_parameters=False):
        self._curr_ckpt_path = os.path.join(save_dir, str(tag if tag is not None else f"global_step{self.global_steps}"))
        return super().save_checkpoint(save_dir, tag, client_state, save_latest, exclude_frozen_parameters)

    def load_checkpoint(self, load_dir, tag=None, **kwargs):
        latest = os.path.join(load_dir, "latest")
        resolved = tag
        if resolved is None and os.path.isfile(latest):
            with open(latest) as fd:
                resolved = fd.read().strip()
        self._curr_ckpt_path = os.path.join(load_dir, str(resolved))
        return super().load_checkpoint(load_dir, tag, **kwargs)
