"""ZeRO stage 3 optimizer (parameter + gradient + optimizer-state sharding).

Thin specialisation of ZeroShardedOptimizer with unit fetch/release hooks and trace-ordered prefetch.
Parity: deepspeed/runtime/zero/stage3.py:73-2756 (+ parameter_offload.py, partitioned_param_coordinator.py).
"""
import torch

from .sharded_optimizer import ZeroShardedOptimizer


class DeepSpeedZeroOptimizer_Stage3(ZeroShardedOptimizer):

    def __init__(self, module, init_optimizer, timers=None, ds_config=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, contiguous_gradients=True, reduce_bucket_size=500000000,
                 prefetch_bucket_size=50000000, max_reuse_distance=1000000000, max_live_parameters=1000000000,
                 param_persistence_threshold=100000, model_persistence_threshold=None, dp_process_group=None, reduce_scatter=True,
                 overlap_comm=False, offload_optimizer_config=None, offload_param_config=None, sub_group_size=1000000000000,
                 mpu=None, clip_grad=0.0, gradient_accumulation_dtype=None, communication_data_type=None,
                 postscale_gradients=True, gradient_predivide_factor=1.0, gradient_accumulation_steps=1, elastic_checkpoint=False,
                 aio_config=None, all2all_process_group=None, zero_hpz_partition_size=1, zero_quantized_weights=False,
                 zero_quantized_nontrainable_weights=False, zero_config=None, dtype=torch.float16, grad_divisor=None,
                 shard_group=None, replica_group=None):
        if zero_config is None:
            from .config import DeepSpeedZeroConfig
            zero_config = DeepSpeedZeroConfig(stage=3, overlap_comm=overlap_comm, offload_optimizer=offload_optimizer_config,
                                              offload_param=offload_param_config, stage3_prefetch_bucket_size=prefetch_bucket_size,
                                              stage3_max_live_parameters=max_live_parameters,
                                              stage3_param_persistence_threshold=param_persistence_threshold,
                                              zero_quantized_weights=zero_quantized_weights)
        super().__init__(module, init_optimizer, stage=3, zero_config=zero_config, dtype=dtype, dp_group=dp_process_group,
                         shard_group=shard_group, replica_group=replica_group, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, clip_grad=clip_grad,
                         gradient_accumulation_steps=gradient_accumulation_steps, grad_accum_dtype=gradient_accumulation_dtype,
                         communication_data_type=communication_data_type, mpu=mpu, timers=timers,
                         gradient_predivide_factor=gradient_predivide_factor, prescale_gradients=not postscale_gradients,
                         grad_divisor=grad_divisor, elastic_checkpoint=elastic_checkpoint, aio_config=aio_config)

    def empty_partition_cache(self):
        """Release every gathered unit (parity: stage3.empty_partition_cache)."""
        for u in self.units:
            self.release_unit(u)


def estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1, cpu_offload=True,
                                          cpu_offload_params=True, zero_init=True, additional_buffer_factor=1.5):
    """Bytes (cpu, gpu, largest_layer) under ZeRO-3 (parity: stage3.py:2613-2650)."""
    total_gpus = num_nodes * num_gpus_per_node
    gpus_factor = 1 / num_nodes
    largest_layer_memory = 4 * largest_layer_params
    if cpu_offload:
        if cpu_offload_params:
            gpu_mem = largest_layer_memory
            cpu_mem = total_params * (18 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 18 * gpus_factor)) * additional_buffer_factor
        else:
            gpu_mem = largest_layer_memory + int(2 * total_params / total_gpus)
            cpu_mem = total_params * (16 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 16 * gpus_factor)) * additional_buffer_factor
    else:
        gpu_mem = largest_layer_memory + int(18 * total_params / total_gpus)
        cpu_mem = largest_layer_params * 4 * num_gpus_per_node * additional_buffer_factor if zero_init else total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem), largest_layer_memory


def model_to_params(model):
    total_params = sum(p.numel() for p in model.parameters())
    largest_layer_params = 0
    for m in model.modules():
        layer_params = sum(p.numel() for p in m.parameters(recurse=False))
        largest_layer_params = max(largest_layer_params, layer_params)
    return total_params, largest_layer_params


def estimate_zero3_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    total_params, largest_layer_params = model_to_params(model)
    estimate_zero3_model_states_mem_needs_all_cold(total_params, largest_layer_params, num_gpus_per_node, num_nodes, additional_buffer_factor)


def estimate_zero3_model_states_mem_needs_all_cold(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    def gb(x):
        return f"{x / 2 ** 30:7.2f}GB"
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} node{'s' if num_nodes > 1 else ''}, {num_gpus_per_node} GPU{'s' if num_gpus_per_node > 1 else ''} per node.\n"
          f"SW: Model with {int(total_params / 1e6)}M total params, {int(largest_layer_params / 1e6)}M largest layer params.")
    print("  per CPU  |  per GPU |   Options")
    for off, offp in ((True, True), (True, False), (False, False)):
        for zi in (True, False):
            cpu, gpu, _ = estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node, num_nodes, off, offp, zi, additional_buffer_factor)
            print(f"  {gb(cpu)} | {gb(gpu)} | offload_param={'cpu' if offp else 'none'}, offload_optimizer={'cpu' if off else 'none'}, zero_init={int(zi)}")
