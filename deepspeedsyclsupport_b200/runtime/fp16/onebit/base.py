"""Shared plumbing of the communication-compressed optimizers: backend selection, error-buffer sizing, the switch that
turns the engine's dense gradient all-reduce off once compression starts, exp_avg_mask handling on checkpoint load.

Under the engine these optimizers run on the flat fp32 master regions of the unified ZeRO store (stage 0), so one
compressed all-reduce covers a whole region instead of one per parameter tensor."""
import numpy as np
import torch

from .... import comm as dist


class CompressedCommOptimizer(torch.optim.Optimizer):

    def _init_comm(self, deepspeed, comm_backend_name, cuda_aware):
        self.deepspeed = deepspeed
        self.comm_backend_name = comm_backend_name
        self.cuda_aware = cuda_aware
        self.using_pipeline = False
        assert dist.is_initialized(), "Please initialize the torch distributed backend."
        mpu = getattr(deepspeed, "mpu", None)
        if comm_backend_name in ("nccl", "hccl"):
            from ...comm.nccl import NcclBackend
            self.using_pipeline = hasattr(self.deepspeed, "pipeline_enable_backward_allreduce")
            self.comm_backend_handle = NcclBackend(mpu)
        elif comm_backend_name == "mpi":
            from ...comm.mpi import MpiBackend
            self.comm_backend_handle = MpiBackend(cuda_aware, mpu)
        else:
            raise ValueError(f"unknown comm_backend_name {comm_backend_name}")
        self.size = self.comm_backend_handle.size
        self.divider = int(self.size * 8 / np.gcd(self.size, 8))

    def _corrected_size(self, n):
        q = self.size * self.divider
        return n if n % q == 0 else n + (q - n % q)

    def _new_error_buffers(self, state, p):
        state["tensor_size"] = p.numel()
        state["corrected_tensor_size"] = self._corrected_size(p.numel())
        state["server_chunk_size"] = state["corrected_tensor_size"] // self.size
        state["worker_error"] = torch.zeros(state["corrected_tensor_size"], device=p.device)
        state["server_error"] = torch.zeros(state["server_chunk_size"], device=p.device)

    def _set_dense_allreduce(self, enabled):
        if self.deepspeed is None:
            return
        if self.using_pipeline:
            self.deepspeed.pipeline_enable_backward_allreduce = enabled
        else:
            self.deepspeed.enable_backward_allreduce = enabled

    def _local_rank(self):
        return getattr(self.deepspeed, "local_rank", 0)

    def _apply_mask(self, group, t):
        if "exp_avg_mask" in group:
            if t.device != group["exp_avg_mask"].device:
                group["exp_avg_mask"] = group["exp_avg_mask"].to(device=t.device)
            t.mul_(group["exp_avg_mask"])

    def _keep_user_masks(self, state_dict):
        # the mask can differ between training phases (e.g. seq 128 vs 512), so the one set by the script wins
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")

    def _drop_error_buffers(self):
        # errors are rank-local and not checkpointed consistently: reset after loading (as upstream does)
        for group in self.param_groups:
            for p in group["params"]:
                self.state[p].pop("worker_error", None)
                self.state[p].pop("server_error", None)

    def _grads(self, group, grads_this_group):
        for i, p in enumerate(group["params"]):
            g = grads_this_group[i] if grads_this_group is not None else None
            if g is None:
                if p.grad is None:
                    continue
                g = p.grad.data
            if g.is_sparse:
                raise RuntimeError(f"{type(self).__name__} does not support sparse gradients")
            yield p, g
