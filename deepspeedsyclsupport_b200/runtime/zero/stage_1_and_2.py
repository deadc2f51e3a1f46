"""ZeRO stage 1 / 2 optimizer (optimizer-state and gradient sharding).

Thin specialisation of ZeroShardedOptimizer: full low-precision unit buffers stay resident; stage 1 reduces the
gradient buffers once at the accumulation boundary, stage 2 reduce-scatters each unit as soon as its backward is
complete and frees the full gradient.  Parity: deepspeed/runtime/zero/stage_1_and_2.py:96-2492.
"""
import torch

from .sharded_optimizer import ZeroShardedOptimizer


class DeepSpeedZeroOptimizer(ZeroShardedOptimizer):

    def __init__(self, init_optimizer, module=None, param_names=None, timers=None, static_loss_scale=1.0,
                 dynamic_loss_scale=False, dynamic_loss_args=None, verbose=True, contiguous_gradients=True,
                 reduce_bucket_size=500000000, use_multi_rank_bucket_allreduce=True, allgather_bucket_size=5000000000,
                 dp_process_group=None, expert_parallel_group=None, expert_data_parallel_group=None, reduce_scatter=True,
                 overlap_comm=False, offload_optimizer_config=None, mpu=None, clip_grad=0.0, gradient_accumulation_dtype=None,
                 communication_data_type=None, postscale_gradients=True, gradient_predivide_factor=1.0,
                 gradient_accumulation_steps=1, ignore_unused_parameters=True, partition_grads=True, round_robin_gradients=False,
                 has_moe_layers=False, fp16_master_weights_and_gradients=False, elastic_checkpoint=False, zero_config=None,
                 dtype=torch.float16, grad_divisor=None, stage=None):
        assert module is not None, "the B200 ZeRO optimizer shards by module structure: pass `module=`"
        if zero_config is None:
            from .config import DeepSpeedZeroConfig
            zero_config = DeepSpeedZeroConfig(stage=2 if partition_grads else 1, overlap_comm=overlap_comm,
                                              offload_optimizer=offload_optimizer_config)
        super().__init__(module, init_optimizer, stage=stage if stage is not None else (2 if partition_grads else 1),
                         zero_config=zero_config, dtype=dtype, dp_group=dp_process_group, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, clip_grad=clip_grad,
                         gradient_accumulation_steps=gradient_accumulation_steps, grad_accum_dtype=gradient_accumulation_dtype,
                         communication_data_type=communication_data_type, mpu=mpu, timers=timers,
                         gradient_predivide_factor=gradient_predivide_factor, prescale_gradients=not postscale_gradients,
                         grad_divisor=grad_divisor, elastic_checkpoint=elastic_checkpoint)


def estimate_zero2_model_states_mem_needs(total_params, num_gpus_per_node=1, num_nodes=1, cpu_offload=True, additional_buffer_factor=1.5):
    """Bytes (cpu, gpu) for params + grads + optimizer states under ZeRO-2 (parity: stage_1_and_2.py:2396-2416)."""
    total_gpus = num_nodes * num_gpus_per_node
    if cpu_offload:
        gpu_mem = 2 * total_params
        cpu_mem = total_params * max(4 * total_gpus, 16) * additional_buffer_factor
    else:
        gpu_mem = 4 * total_params + int(16 * total_params / total_gpus)
        cpu_mem = total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem)


def _fmt_options(fn, total_params, largest=None, **kw):
    def gb(x):
        return f"{x / 2 ** 30:7.2f}GB"
    lines = []
    for offload in (True, False):
        cpu, gpu = fn(total_params, cpu_offload=offload, **kw)
        lines.append(f"  {gb(cpu)} | {gb(gpu)} | offload_optimizer={'cpu' if offload else 'none'}")
    return lines


def estimate_zero2_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    total_params = sum(p.numel() for p in model.parameters())
    return estimate_zero2_model_states_mem_needs_all_cold(total_params, num_gpus_per_node, num_nodes, additional_buffer_factor)


def estimate_zero2_model_states_mem_needs_all_cold(total_params, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} node{'s' if num_nodes > 1 else ''}, {num_gpus_per_node} GPU{'s' if num_gpus_per_node > 1 else ''} per node.\n"
          f"SW: Model with {int(total_params / 1e6)}M total params.")
    print("  per CPU  |  per GPU |   Options")
    for l in _fmt_options(estimate_zero2_model_states_mem_needs, total_params, num_gpus_per_node=num_gpus_per_node,
                          num_nodes=num_nodes, additional_buffer_factor=additional_buffer_factor):
        print(l)
