"""PipelineEngine: executes a 1F1B TrainSchedule / InferenceSchedule instruction stream over NCCL p2p.

Parity: deepspeed/runtime/pipe/engine.py:55-1372 (train_batch :321, eval_batch, _exec_* handlers :651-1204,
_INSTRUCTION_MAP :1344, _exec_schedule :1357, meta handshake :854-970, tied-weight / DP gradient reduction :257-289,
loss aggregation :563-589).  Pipeline-stage transfers stay on NCCL send/recv by design (BASELINE north star); ZeRO
stages 0/1 compose with PP (assert :71-72 of the reference).
"""
from types import MethodType

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


def is_even(number):
    return number % 2 == 0


class PipelineEngine(DeepSpeedEngine):
    ID_TO_DTYPE = [torch.float32, torch.float64, torch.complex64, torch.complex128, torch.float16, torch.bfloat16, torch.uint8,
                   torch.int8, torch.int16, torch.int32, torch.int64, torch.bool]
    DTYPE_TO_ID = {dtype: id_ for id_, dtype in enumerate(ID_TO_DTYPE)}

    def __init__(self, has_bool_tensors=False, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        self.enable_backward_allreduce = False
        self.has_bool_tensors = has_bool_tensors
        self.eval_return_logits = False
        self.outputs = None
        self.using_bf16_optimizer = False
        self.pipeline_enable_backward_allreduce = True
        self.log_batch_step_id = -1
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        self.grid = self.module._grid
        self.global_rank = self.grid.get_global_rank()
        assert self.dp_world_size == self.grid.data_parallel_size
        assert self.train_batch_size() == self.micro_batch_size * self.micro_batches * self.grid.data_parallel_size
        self.num_stages = self.grid.pipe_parallel_size
        self.stage_id = self.grid.get_stage_id()
        self.prev_stage = self.stage_id - 1
        self.next_stage = self.stage_id + 1
        self.data_iterator = None
        self.batch_fn = None
        self._force_grad_boundary = False
        self.batch_timer = None
        self.is_pipe_parallel = self.grid.pipe_parallel_size > 1
        self.is_data_parallel = self.grid.data_parallel_size > 1
        self.is_model_parallel = self.grid.model_parallel_size > 1
        self.is_pipe_partitioned = self.is_model_parallel and self._config.pipeline["pipe_partitioned"]
        self.is_grad_partitioned = self.is_model_parallel and self._config.pipeline["grad_partitioned"]
        self.module.checkpoint_parallel_write_pipeline = self._config.checkpoint_parallel_write_pipeline
        if self.training_data:
            self._build_data_iter(self.training_data)
        self.num_pipe_buffers = 0
        self.pipe_buffers = {"inputs": [], "labels": [], "outputs": [], "output_tensors": []}
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None
        self.first_output_send = True
        self.first_gradient_send = True
        self.pipe_partition_input_meta_cache = None
        self.pipe_partition_output_meta_cache = None
        self.pipe_partition_grad_meta_cache = None
        self.grad_partition_grad_layer_meta_cache = None
        self.loss = torch.tensor(0.0).to(self.device)
        self.total_loss = None
        self.agg_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        self.dp_group_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        if self._config.pipeline["activation_checkpoint_interval"] > 0:
            self.module.activation_checkpoint_interval = self._config.pipeline["activation_checkpoint_interval"]
        self.module.__class__.loss_fn = self.module.loss_fn if not hasattr(self.module.__class__, "loss_fn") else self.module.__class__.loss_fn
        self.has_attention_mask = self.module.__class__.__name__ == "GPT2ModelPipe"
        if self.is_pipe_parallel:
            p2p.init_process_groups(self.grid)
        self._last_stage = self.stage_id == self.num_stages - 1
        self._first_stage = self.stage_id == 0

    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def _build_data_iter(self, dataset):
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        pipe_dataloader = self.deepspeed_io(dataset, data_sampler=sampler)
        pipe_dataloader = RepeatingLoader(pipe_dataloader)
        self.set_dataloader(pipe_dataloader)

    def _exec_reduce_tied_grads(self):
        # tied-weight gradients live in the ZeRO gradient buffers: all-reduce the views across the tie group
        for weight, group in self.module.get_tied_weights_and_groups():
            opt = getattr(weight, "ds_zero_optimizer", None)
            if opt is not None:
                grad = opt.main_grad_view(weight)
            else:
                grad = weight.grad
            if grad is not None:
                dist.all_reduce(grad, group=group)

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        if self.pipeline_enable_backward_allreduce:
            self.optimizer.end_micro_step(True)
        self._force_grad_boundary = False

    def _reserve_pipe_buffers(self, num_buffers):
        if self.num_pipe_buffers >= num_buffers:
            return
        num_added = num_buffers - self.num_pipe_buffers
        for key in self.pipe_buffers:
            self.pipe_buffers[key].extend([None] * num_added)
        self.num_pipe_buffers = num_buffers

    def reset_activation_shape(self):
        self.first_output_send = True
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None
        self.pipe_partition_input_meta_cache = None
        self.pipe_partition_output_meta_cache = None
        self.pipe_partition_grad_meta_cache = None
        self.grad_partition_grad_layer_meta_cache = None

    def train_batch(self, data_iter=None):
        """One full global batch: micro_batches x (fwd, bwd) through the pipeline, gradient reduction, optimizer step.
        Returns the mean loss over the batch (valid on every rank)."""
        if not torch._C.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss = None
        self._compute_loss = True
        sched = schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        self._exec_schedule(sched)
        self.agg_train_loss = self._aggregate_total_loss()
        if self.global_steps % self.steps_per_print() == 0 and self.global_rank == 0:
            print(f"steps: {self.global_steps} loss: {self.agg_train_loss:0.4f} ")
        if self.monitor.enabled and self.global_rank == 0:
            self.monitor.write_events([("Train/Samples/train_loss", self.agg_train_loss.mean().item(), self.global_samples)])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, compute_loss=True, reduce_output="avg", bcast_loss=True, num_micro_batches=None):
        self.eval_return_logits = return_logits
        self.module.eval()
        self._compute_loss = compute_loss
        train_iterator = self.data_iterator
        self.set_dataiterator(data_iter)
        micro_batches = self.micro_batches if num_micro_batches is None else num_micro_batches
        sched = schedule.InferenceSchedule(micro_batches=micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        dist.barrier()
        with torch.no_grad():
            self._exec_schedule(sched)
        eval_output = None
        if self._last_stage:
            eval_output = self._reduce_outputs(self.fwd_outputs, reduce=reduce_output, micro_batches=micro_batches)
        if compute_loss and (bcast_loss or self.monitor.enabled):
            eval_output = self._bcast_pipe_scalar(eval_output)
        self.set_dataiterator(train_iterator)
        if return_logits:
            outputs = self.outputs
            self.outputs = None
            return eval_output, outputs
        return eval_output

    def set_train_batch_size(self, train_batch_size):
        super().set_train_batch_size(train_batch_size)
        self.micro_batches = self.gradient_accumulation_steps()

    def is_first_stage(self):
        """True if this process is in the first stage of the pipeline (reference: pipe/engine.py:509)."""
        return self._first_stage

    def is_last_stage(self):
        return self._last_stage

    def log_for_device(self, *msg):
        """Print from every pipeline stage / data-parallel replica selected by DSB200_PIPE_LOG_STAGE / _DP (-1 = all)."""
        import os
        want_stage, want_dp = int(os.environ.get("DSB200_PIPE_LOG_STAGE", "-2")), int(os.environ.get("DSB200_PIPE_LOG_DP", "-1"))
        if want_stage in (-1, self.stage_id) and want_dp in (-1, self.grid.data_parallel_id):
            print(f"RANK={dist.get_rank()} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} STEP-ID={self.global_steps}", *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.global_steps % self.steps_per_print() == 0:
            print(*msg)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        """Allocated / cached device memory and their change since the previous call, on one rank."""
        acc = get_accelerator()
        if not acc.is_available() or (print_rank != -1 and self.global_rank != print_rank):
            return
        acc.synchronize()
        if reset_max:
            acc.reset_max_memory_allocated()
        alloc, cached = acc.memory_allocated(), acc.memory_reserved() if hasattr(acc, "memory_reserved") else 0
        prev = getattr(self, "_mem_status_prev", (0, 0))
        self._mem_status_prev = (alloc, cached)
        gib = 1024**3
        print(f"RANK={self.global_rank} STAGE={self.stage_id} STEP={self.global_steps} MEMSTATS {msg} "
              f"current alloc={alloc / gib:0.4f}GB (delta={(alloc - prev[0]) / gib:0.4f}GB max={acc.max_memory_allocated() / gib:0.4f}GB) "
              f"current cache={cached / gib:0.4f}GB (delta={(cached - prev[1]) / gib:0.4f}GB)", flush=True)

    def is_first_stage_fn(self):
        return self.stage_id == 0

    def is_last_stage_fn(self):
        return self.stage_id == self.num_stages - 1

    def _reduce_outputs(self, outputs, reduce="avg", reduce_dp=True, micro_batches=None):
        if reduce is None:
            return outputs
        if reduce.lower() == "avg":
            if torch.is_tensor(outputs[0]):
                reduced = sum(outputs)
            else:
                assert isinstance(outputs, (list, tuple))
                reduced = [torch.zeros_like(o) for o in outputs[0]]
                for idx, out in outputs:
                    reduced[idx] += out
            if torch.is_tensor(reduced):
                reduced /= micro_batches if micro_batches else self.micro_batches
            if reduce_dp and self.is_data_parallel:
                if torch.is_tensor(reduced):
                    dist.all_reduce(reduced, group=self.mpu.get_data_parallel_group())
                    reduced /= self.dp_world_size
            return reduced
        raise NotImplementedError(f"reduction type {reduce} not supported.")

    def _bcast_pipe_scalar(self, data, src_rank=None, dtype=torch.float32):
        if src_rank is None:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
        assert src_rank in self.grid.pp_group
        if self.global_rank == src_rank:
            result = data.clone().detach().type(dtype).to(self.device)
        else:
            result = torch.Tensor([0.]).type(dtype).to(self.device)
        dist.broadcast(tensor=result, src=src_rank, group=self.mpu.get_pipe_parallel_group())
        return result

    def _aggregate_total_loss(self):
        if self._last_stage:
            loss = self._scale_loss_by_gas(self.total_loss)
            self.dp_group_loss = loss.clone().detach()
            agg_loss = self.dp_group_loss.clone().detach()
            if self.is_data_parallel:
                dist.all_reduce(agg_loss, group=self.mpu.get_data_parallel_group())
                agg_loss /= self.dp_world_size
            assert self.global_rank in self.grid.pp_group
            losses = torch.stack([self.dp_group_loss, agg_loss]).float()
            if self.is_pipe_parallel:
                dist.broadcast(tensor=losses, src=self.global_rank, group=self.mpu.get_pipe_parallel_group())
        else:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
            assert src_rank in self.grid.pp_group
            losses = torch.Tensor([0., 0.]).to(self.device)
            dist.broadcast(tensor=losses, src=src_rank, group=self.grid.get_pipe_parallel_group())
            self.dp_group_loss = losses[0].clone().detach()
            agg_loss = losses[1].clone().detach()
        return agg_loss

    def set_dataloader(self, loader):
        if self._first_stage or self._last_stage:
            self.training_dataloader = loader
            self.data_iterator = iter(self.training_dataloader)

    def set_dataiterator(self, iterator):
        if self._first_stage or self._last_stage:
            self.training_dataloader = None
            self.data_iterator = iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def is_gradient_accumulation_boundary(self):
        return self._force_grad_boundary

    def _next_batch(self):
        batch = None
        if self.data_iterator is not None:
            batch = next(self.data_iterator)
        if self.batch_fn:
            batch = self.batch_fn(batch)
        return batch

    def _exec_forward_pass(self, buffer_id):
        self.tput_timer.start()
        if isinstance(self.pipe_buffers["inputs"][buffer_id], tuple):
            inputs = tuple(t.clone() for t in self.pipe_buffers["inputs"][buffer_id])
        else:
            inputs = self.pipe_buffers["inputs"][buffer_id].clone()
        self._zero_grads(inputs)
        outputs = self.module.forward(inputs) if not isinstance(inputs, tuple) or len(inputs) > 1 else self.module.forward(inputs[0])
        self.pipe_buffers["outputs"][buffer_id] = outputs
        if self._last_stage:
            if self._compute_loss and self.module.loss_fn is not None:
                labels = self.pipe_buffers["labels"][buffer_id]
                self.loss = self.module.loss_fn(outputs, labels)
            else:
                self.loss = outputs
            if self.eval_return_logits:
                self.outputs = outputs
            if isinstance(self.loss, torch.Tensor):
                self.fwd_outputs.append(self.loss.detach())
                if self.total_loss is None:
                    self.total_loss = torch.zeros_like(self.loss)
                self.total_loss += self.loss.detach()
            else:
                self.fwd_outputs.append([l.detach() for l in self.loss])
                if self.total_loss is None:
                    self.total_loss = [torch.zeros_like(l) for l in self.loss]
                for idx, l in enumerate(self.loss):
                    self.total_loss[idx] += l.detach()

    def _exec_backward_pass(self, buffer_id):
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        if self._last_stage:
            loss = self._scale_loss_by_gas(self.loss.float()) if self.gradient_accumulation_steps() > 1 else self.loss
            self.optimizer.backward(loss)
            self.optimizer.micro_step_id += 1
            return
        outputs = self.pipe_buffers["outputs"][buffer_id]
        grad_tensors = self.grad_layer
        if isinstance(outputs, tuple):
            out_tensors = [t for t in outputs if t.is_floating_point()]
            assert len(out_tensors) == len(grad_tensors)
            torch.autograd.backward(tensors=out_tensors, grad_tensors=grad_tensors)
        else:
            torch.autograd.backward(tensors=(outputs, ), grad_tensors=(grad_tensors, ))
        self.pipe_buffers["output_tensors"][buffer_id] = None
        self.pipe_buffers["outputs"][buffer_id] = None
        grad_tensors = None

    def _exec_load_micro_batch(self, buffer_id):
        batch = self._next_batch()
        if self._first_stage:
            loaded = None
            if torch.is_tensor(batch[0]):
                loaded = batch[0].clone().to(self.device).detach()
                loaded.requires_grad = loaded.is_floating_point()
            else:
                assert isinstance(batch[0], (tuple, list))
                loaded = []
                for x in batch[0]:
                    assert torch.is_tensor(x)
                    mine = x.clone().detach().to(self.device)
                    mine.requires_grad = mine.is_floating_point()
                    loaded.append(mine)
                loaded = tuple(loaded)
            self.pipe_buffers["inputs"][buffer_id] = loaded
        if self._last_stage:
            loaded = batch[1]
            if torch.is_tensor(batch[1]):
                loaded = batch[1].to(self.device)
            elif isinstance(batch[1], (tuple, list)):
                loaded = tuple(x.to(self.device).detach() for x in batch[1])
            self.pipe_buffers["labels"][buffer_id] = loaded

    # ---- meta handshake: (type, ndim, shape, dtype) so the receiver can allocate buffers once
    def _send_tensor_meta(self, buffer, recv_stage):
        if isinstance(buffer, torch.Tensor):
            type_tensor = torch.LongTensor(data=[0]).to(self.device)
            p2p.send(type_tensor, recv_stage)
            send_shape = torch.LongTensor(data=list(buffer.size())).to(self.device)
            send_ndims = torch.LongTensor(data=[len(buffer.size())]).to(self.device)
            send_dtype = torch.LongTensor(data=[self.DTYPE_TO_ID[buffer.dtype]]).to(self.device)
            p2p.send(send_ndims, recv_stage)
            p2p.send(send_shape, recv_stage)
            p2p.send(send_dtype, recv_stage)
        elif isinstance(buffer, (tuple, list)):
            type_tensor = torch.LongTensor(data=[2 if isinstance(buffer, tuple) else 1]).to(self.device)
            p2p.send(type_tensor, recv_stage)
            count_tensor = torch.LongTensor(data=[len(buffer)]).to(self.device)
            p2p.send(count_tensor, recv_stage)
            for tensor in buffer:
                assert isinstance(tensor, torch.Tensor)
                send_shape = torch.LongTensor(data=list(tensor.size())).to(self.device)
                send_ndims = torch.LongTensor(data=[len(tensor.size())]).to(self.device)
                send_dtype = torch.LongTensor(data=[self.DTYPE_TO_ID[tensor.dtype]]).to(self.device)
                p2p.send(send_dtype, recv_stage)
                p2p.send(send_ndims, recv_stage)
                p2p.send(send_shape, recv_stage)
        else:
            raise NotImplementedError(f"Could not send meta type {type(buffer)}")

    def _recv_tensor_meta(self, send_stage):
        type_tensor = torch.LongTensor(data=[0]).to(self.device)
        p2p.recv(type_tensor, send_stage)
        recv_type = type_tensor.item()
        if recv_type == 0:
            recv_ndims = torch.LongTensor(data=[0]).to(self.device)
            p2p.recv(recv_ndims, send_stage)
            recv_shape = torch.LongTensor([1] * recv_ndims.item()).to(self.device)
            p2p.recv(recv_shape, send_stage)
            recv_dtype = torch.LongTensor(data=[0]).to(self.device)
            p2p.recv(recv_dtype, send_stage)
            return self._allocate_buffer(recv_shape.tolist(), dtype=self.ID_TO_DTYPE[recv_dtype.item()], num_buffers=1)[0]
        elif recv_type in (1, 2):
            count_tensor = torch.LongTensor(data=[0]).to(self.device)
            p2p.recv(count_tensor, send_stage)
            num_tensors = count_tensor.item()
            buffers = []
            for idx in range(num_tensors):
                recv_dtype = torch.LongTensor(data=[0]).to(self.device)
                p2p.recv(recv_dtype, send_stage)
                recv_ndims = torch.LongTensor(data=[0]).to(self.device)
                p2p.recv(recv_ndims, send_stage)
                recv_shape = torch.LongTensor([1] * recv_ndims.item()).to(self.device)
                p2p.recv(recv_shape, send_stage)
                buffers.append(self._allocate_buffer(recv_shape.tolist(), dtype=self.ID_TO_DTYPE[recv_dtype.item()], num_buffers=1)[0])
            if recv_type == 2:
                buffers = tuple(buffers)
            return buffers
        raise NotImplementedError(f"Could not receive type {type(recv_type)}")

    def _exec_send_activations(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.first_output_send:
            self.first_output_send = False
            self._send_tensor_meta(outputs, self.next_stage)
        if isinstance(outputs, torch.Tensor):
            p2p.send(outputs, self.next_stage)
        elif isinstance(outputs, tuple):
            for idx, buffer in enumerate(outputs):
                p2p.send(buffer, self.next_stage)
        else:
            raise NotImplementedError(f"Could not send output of type {type(outputs)}")

    def _exec_send_grads(self, buffer_id):
        inputs = self.pipe_buffers["inputs"][buffer_id]
        if isinstance(inputs, torch.Tensor):
            assert inputs.grad is not None
            p2p.send(inputs.grad, self.prev_stage)
        else:
            for idx, buffer in enumerate(inputs):
                if not buffer.is_floating_point():
                    assert buffer.grad is None
                    continue
                assert buffer.grad is not None
                p2p.send(buffer.grad, self.prev_stage)
        self.pipe_buffers["inputs"][buffer_id] = None

    def _exec_recv_activations(self, buffer_id):
        recvd = None
        if self.pipe_recv_buf is None:
            self.pipe_recv_buf = self._recv_tensor_meta(self.prev_stage)
        if isinstance(self.pipe_recv_buf, torch.Tensor):
            p2p.recv(self.pipe_recv_buf, self.prev_stage)
            recvd = self.pipe_recv_buf.clone().detach()
            recvd.requires_grad = recvd.is_floating_point()
        else:
            assert isinstance(self.pipe_recv_buf, tuple)
            recvd = [None] * len(self.pipe_recv_buf)
            for idx, buffer in enumerate(self.pipe_recv_buf):
                assert torch.is_tensor(buffer)
                p2p.recv(buffer, self.prev_stage)
                recvd[idx] = buffer.clone().detach()
            recvd = tuple(recvd)
            for buffer in recvd:
                buffer.requires_grad = buffer.is_floating_point()
        self.pipe_buffers["inputs"][buffer_id] = recvd

    def _exec_recv_grads(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.grad_layer is None:
            if isinstance(outputs, torch.Tensor):
                s = list(outputs.size())
                self.grad_layer = self._allocate_buffer(s, dtype=outputs.dtype, num_buffers=1)[0]
            else:
                sizes_and_dtypes = [(list(t.size()), t.dtype) for t in outputs if t.is_floating_point()]
                self.grad_layer = self._allocate_buffers(sizes_and_dtypes, num_buffers=1)[0]
        if isinstance(self.grad_layer, torch.Tensor):
            p2p.recv(self.grad_layer, self.next_stage)
        else:
            assert isinstance(outputs, tuple)
            for idx, buffer in enumerate(self.grad_layer):
                p2p.recv(buffer, self.next_stage)

    def _exec_optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        self.micro_steps += self.micro_batches

    def _zero_grads(self, inputs):
        if isinstance(inputs, torch.Tensor):
            if inputs.grad is not None:
                inputs.grad.data.zero_()
        else:
            for t in inputs:
                if t.grad is not None:
                    t.grad.data.zero_()

    def _allocate_zeros(self, shape, **kwargs):
        if "dtype" not in kwargs:
            if self.fp16_enabled():
                kwargs["dtype"] = torch.half
            if self.bfloat16_enabled():
                kwargs["dtype"] = torch.bfloat16
        return torch.zeros(shape, device=self.device, **kwargs)

    def _allocate_buffer(self, shape, num_buffers=-1, **kwargs):
        buffers = []
        if num_buffers == -1:
            num_buffers = self.num_pipe_buffers
        for count in range(num_buffers):
            buffers.append(self._allocate_zeros(shape, **kwargs))
        return buffers

    def _allocate_buffers(self, shapes_and_dtypes, requires_grad=False, num_buffers=-1):
        buffers = []
        if num_buffers == -1:
            num_buffers = self.num_pipe_buffers
        for count in range(num_buffers):
            buffer = []
            for shape, dtype in shapes_and_dtypes:
                buffer.append(self._allocate_zeros(shape, dtype=dtype, requires_grad=requires_grad))
            buffers.append(buffer)
        return buffers

    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def module_state_dict(self, exclude_frozen_parameters=False):
        """Layers are saved to their own files by PipelineModule.save_state_dict (layer_XX-model_YY-model_states.pt)."""
        assert isinstance(self.module, PipelineModule)
        assert self._curr_ckpt_path is not None, "PipelineEngine expects module_state_dict() to be called from save_checkpoint()"
        self.module.save_state_dict(self._curr_ckpt_path, checkpoint_engine=self.checkpoint_engine,
                                    exclude_frozen_params=exclude_frozen_parameters)
        return None

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False, state_dict=None):
        assert custom_load_fn is None, "custom_load_fn not supported w. pipeline parallelism"
        state_dict = checkpoint["module"] if checkpoint is not None else state_dict
        if (state_dict is not None) and (not isinstance(state_dict, str)):
            super().load_module_state_dict(state_dict=state_dict, strict=strict)
            return
        self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict, checkpoint_engine=self.checkpoint_engine)
        if self.optimizer is not None and hasattr(self.optimizer, "refresh_from_params"):
            self.optimizer.refresh_from_params()

    _curr_ckpt_path = None

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True, exclude_frozen_parameters=False):
        import os
        t = tag if tag is not None else f"global_step{self.global_steps}"
        self._curr_ckpt_path = os.path.join(save_dir, str(t))
        return super().save_checkpoint(save_dir, tag, client_state, save_latest, exclude_frozen_parameters)

    def load_checkpoint(self, load_dir, tag=None, **kwargs):
        import os
        t = tag
        if t is None and os.path.isfile(os.path.join(load_dir, "latest")):
            t = open(os.path.join(load_dir, "latest")).read().strip()
        self._curr_ckpt_path = os.path.join(load_dir, str(t))
        return super().load_checkpoint(load_dir, tag, **kwargs)

    _INSTRUCTION_MAP = {
        schedule.OptimizerStep: _exec_optimizer_step,
        schedule.ReduceGrads: _exec_reduce_grads,
        schedule.ReduceTiedGrads: _exec_reduce_tied_grads,
        schedule.LoadMicroBatch: _exec_load_micro_batch,
        schedule.ForwardPass: _exec_forward_pass,
        schedule.BackwardPass: _exec_backward_pass,
        schedule.SendActivation: _exec_send_activations,
        schedule.RecvActivation: _exec_recv_activations,
        schedule.SendGrad: _exec_send_grads,
        schedule.RecvGrad: _exec_recv_grads,
    }

    def _exec_schedule(self, pipe_schedule):
        self._reserve_pipe_buffers(pipe_schedule.num_pipe_buffers())
        self.fwd_outputs = []
        for step_cmds in pipe_schedule:
            for cmd in step_cmds:
                if type(cmd) not in self._INSTRUCTION_MAP:
                    raise RuntimeError(f"{self.__class__.__name__} does not understand instruction {repr(cmd)}")
                self._exec_instr = MethodType(self._INSTRUCTION_MAP[type(cmd)], self)
                self._exec_instr(**cmd.kwargs)
