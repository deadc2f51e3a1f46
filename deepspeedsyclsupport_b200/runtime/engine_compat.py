"""Engine surface that user code and Megatron-style integrations call but the flat-store design does not need internally:
the remaining config getters of the reference engine and its "fallback" gradient all-reduce family (dense buckets and
sparse embedding gradients over `param.grad`), kept functional so that code written against the reference keeps working.
(reference: deepspeed/runtime/engine.py:473-879 getters, :2326-2518 allreduce_bucket / buffered_allreduce_fallback /
sparse_allreduce*, :2952 update_optimizer_step)"""
import torch

from .. import comm as dist
from ..utils.logging import log_dist
from .sparse_tensor import SparseTensor


class EngineCompatMixin(object):

    # ------------------------------------------------------------------ config getters
    def eigenvalue_verbose(self):
        return self._config.eigenvalue_verbose

    def eigenvalue_max_iter(self):
        return self._config.eigenvalue_max_iter

    def eigenvalue_tol(self):
        return self._config.eigenvalue_tol

    def eigenvalue_stability(self):
        return self._config.eigenvalue_stability

    def eigenvalue_gas_boundary_resolution(self):
        return self._config.eigenvalue_gas_boundary_resolution

    def eigenvalue_layer_name(self):
        return self._config.eigenvalue_layer_name

    def eigenvalue_layer_num(self):
        return self._config.eigenvalue_layer_num

    def data_sampling_config(self):
        return self._config.data_efficiency_config.get("data_sampling", {})

    def curriculum_learning_config(self):
        return self.data_sampling_config().get("curriculum_learning", {})

    def quantize_training(self):
        sp = self._config.compression_config.get("weight_quantization", {}).get("shared_parameters", {})
        return (sp.get("quantize_weight_in_forward", False), sp.get("enabled", False), sp.get("quantize_groups", 1),
                sp.get("fp16_mixed_quantize", {}).get("enabled", False) if isinstance(sp.get("fp16_mixed_quantize"), dict)
                else sp.get("fp16_mixed_quantize", False),
                sp.get("quantize_change_ratio", 0.001), sp.get("quantization_type", "symmetric"), sp.get("rounding", "nearest"),
                sp.get("quantize_verbose", False), sp.get("quantizer_kernel", False))

    def zero_partial_offload(self):
        return getattr(self._config.zero_config.offload_optimizer, "ratio", 1.0)

    def zero_multi_rank_bucket_allreduce(self):
        return self._config.zero_config.use_multi_rank_bucket_allreduce

    def zero_allgather_partitions(self):
        return self._config.zero_config.allgather_partitions

    def is_first_weights_partition_group(self):
        mics = self._config.zero_config.mics_shard_size
        if mics > 0:
            return self.global_rank < mics
        return self.zero_optimization_stage() == 3

    def zero_has_nvme_offload(self):
        opt = self.optimizer
        return bool(getattr(opt, "swap_optimizer", False) or getattr(opt, "params_in_nvme_and_cpu", False))

    def dump_state(self):
        return self._config.dump_state

    def is_elastic_model_parallel_supported(self):
        if not self.elasticity_enabled():
            return False
        from ..elasticity.utils import is_torch_elastic_compatible  # noqa: F401
        return self._config.elastic_model_parallel_size >= 1 and self._config.num_gpus_per_node >= 1

    def set_data_post_process_func(self, post_process_func):
        if self.training_dataloader is not None and hasattr(self.training_dataloader, "post_process_func"):
            self.training_dataloader.post_process_func = post_process_func

    @staticmethod
    def is_map_style_dataset(obj):
        return hasattr(obj, "__getitem__") and hasattr(obj, "__len__")

    @staticmethod
    def is_iterable_style_dataset(obj):
        return isinstance(obj, torch.utils.data.IterableDataset)

    def print_forward_breakdown(self, fwd_time):
        """MoE gate / experts / all-to-all share of a forward (reference engine.py:1883-1915)."""
        gate = moe = falltoall = salltoall = 0.0
        for m in self.module.modules():
            if hasattr(m, "time_falltoall"):  # MOELayer
                moe += getattr(m, "time_moe", 0.0)
                falltoall += m.time_falltoall
                salltoall += getattr(m, "time_salltoall", 0.0)
            if hasattr(m, "gate_time"):  # TopKGate
                gate += m.gate_time
        log_dist(f"time (ms) | fwd: {fwd_time:.2f} (fwd_moe: {moe:.2f}, 1st_a2a: {falltoall:.2f}, 2nd_a2a: {salltoall:.2f}, "
                 f"top_k: {gate:.2f})", ranks=[0])

    def all_gather_scalar(self, value, dp_group):
        out = [value.new_zeros(value.size()) for _ in range(dist.get_world_size(group=dp_group))]
        dist.all_gather(out, value, group=dp_group)
        return out

    @staticmethod
    def load_moe_state_dict(checkpoint_path, tag, state_dict, old_moe_load, model, mpu=None, num_experts=1, checkpoint_engine=None):
        from ..moe.utils import load_moe_expert_files
        return load_moe_expert_files(checkpoint_path, tag, state_dict, old_moe_load, model, mpu, num_experts, checkpoint_engine)

    def update_optimizer_step(self, step):
        """Force the step counter of every parameter group / state entry (used after loading a universal checkpoint)."""

        def put(d):
            d["step"] = torch.tensor(step, dtype=d["step"].dtype, device=d["step"].device) if torch.is_tensor(d.get("step")) else step

        opt = self.optimizer
        base = getattr(opt, "optimizer", opt)
        for group in opt.param_groups:
            if "step" in group:
                put(group)
            for p in group["params"]:
                st = base.state.get(p) if hasattr(base, "state") else None
                if st is not None and "step" in st:
                    put(st)
        if hasattr(opt, "set_step"):
            opt.set_step(step)

    # ------------------------------------------------------------------ gradient all-reduce over param.grad
    def _dp_group(self):
        return self.seq_data_parallel_group if getattr(self, "seq_data_parallel_group", None) is not None else self.data_parallel_group

    def allreduce_bucket(self, bucket, dp_group=None):
        """One flat all-reduce of a list of same-dtype tensors; returns the reduced flat tensor (averaged over the group,
        with the configured pre/post-divide split)."""
        group = dp_group if dp_group is not None else self._dp_group()
        world = dist.get_world_size(group=group)
        flat = torch._utils._flatten_dense_tensors(bucket)
        if self.communication_data_type is not None and flat.dtype != self.communication_data_type:
            flat = flat.to(self.communication_data_type)
        if self.postscale_gradients():
            pre = self.gradient_predivide_factor()
            if pre != 1.0:
                flat.mul_(1.0 / pre)
            dist.all_reduce(flat, group=group)
            if self.gradient_average and pre != world:
                flat.mul_(pre / world)
        else:
            flat.mul_(1.0 / world)
            dist.all_reduce(flat, group=group)
        return flat.to(bucket[0].dtype) if flat.dtype != bucket[0].dtype else flat

    def allreduce_and_copy(self, small_bucket, dp_group=None):
        reduced = self.allreduce_bucket(small_bucket, dp_group)
        for buf, synced in zip(small_bucket, torch._utils._unflatten_dense_tensors(reduced, small_bucket)):
            buf.copy_(synced)

    def allreduce_no_retain(self, bucket, dp_group=None, numel_per_bucket=500000000):
        small, numel = [], 0
        for t in bucket:
            small.append(t)
            numel += t.numel()
            if numel > numel_per_bucket:
                self.allreduce_and_copy(small, dp_group)
                small, numel = [], 0
        if small:
            self.allreduce_and_copy(small, dp_group)

    def _grads_by_kind(self, grads=None):
        """-> ({dtype: [dense grads]}, [SparseTensor]) over the module's parameters (missing grads become zeros so every
        rank contributes the same buckets)."""
        dense, sparse = {}, []
        if grads is None:
            grads = []
            for name, p in self.module.named_parameters():
                if not p.requires_grad:
                    continue
                if p.grad is None:
                    p.grad = torch.zeros(p.size(), dtype=p.dtype, device=p.device)
                grads.append((name, p.grad))
        for name, g in grads:
            if g.is_sparse or (self.sparse_gradients_enabled() and name.rsplit(".", 1)[0] in getattr(self, "sparse_tensor_module_names", ())):
                sparse.append(SparseTensor(g.data if not g.is_sparse else g))
            else:
                dense.setdefault(g.dtype, []).append(g.data)
        return dense, sparse

    def buffered_allreduce_fallback(self, grads=None, elements_per_buffer=500000000):
        """Average `param.grad` of every trainable parameter over the data-parallel group: dense gradients in flat buckets,
        sparse (embedding) gradients as (indices, values) all-gathers."""
        dense, sparse = self._grads_by_kind(grads)
        for bucket in dense.values():
            self.allreduce_no_retain(bucket, numel_per_bucket=elements_per_buffer)
        if sparse:
            self.sparse_allreduce_no_retain(sparse)

    def sparse_allreduce_no_retain(self, bucket, dp_group=None):
        for st in self.sparse_allreduce_bucket(bucket, dp_group):
            if st.orig_dense_tensor is not None:
                dense = st.to_dense()
                if st.orig_dense_tensor.is_sparse:
                    st.orig_dense_tensor.data = dense.to_sparse()
                else:
                    st.orig_dense_tensor.copy_(dense)

    def sparse_allreduce_bucket(self, bucket, dp_group=None):
        return [self.sparse_allreduce(st, dp_group) for st in bucket]

    def sparse_allreduce(self, sparse, dp_group=None):
        """Sum of a row-sparse gradient over the group without densifying: every rank contributes its (row index, row) pairs,
        the concatenation is the sum in COO form (duplicates add up in to_dense / coalesce)."""
        group = dp_group if dp_group is not None else self._dp_group()
        world = dist.get_world_size(group=group)
        values = sparse.values
        if self.postscale_gradients():
            pre = self.gradient_predivide_factor()
            if pre != 1.0:
                values = values * (1.0 / pre)
            post = (pre / world) if self.gradient_average else 1.0
        else:
            values, post = values * (1.0 / world), 1.0
        idx = self.sparse_all_gather(sparse.indices, group)
        val = self.sparse_all_gather(values, group)
        sparse.indices = torch.cat(idx)
        sparse.values = torch.cat(val)
        if post != 1.0:
            sparse.values = sparse.values * post
        return sparse

    def sparse_all_gather(self, value, dp_group=None):
        """all-gather of tensors whose first dimension differs per rank (pad to the largest, trim after)."""
        group = dp_group if dp_group is not None else self._dp_group()
        world = dist.get_world_size(group=group)
        n = torch.tensor([value.shape[0]], dtype=torch.long, device=value.device)
        sizes = self.all_gather_scalar(n, group)
        longest = int(max(int(s.item()) for s in sizes))
        pad = value.new_zeros((longest - value.shape[0], ) + tuple(value.shape[1:]))
        mine = torch.cat([value, pad]) if pad.numel() or longest != value.shape[0] else value
        out = [mine.new_empty(mine.shape) for _ in range(world)]
        dist.all_gather(out, mine.contiguous(), group=group)
        return [o[:int(s.item())] for o, s in zip(out, sizes)]
