"""InferenceEngine (v1): dtype conversion, tensor-parallel group creation, AutoTP / injection-policy module
replacement, optional CUDA-graph capture of the forward, generate() passthrough.
Parity: deepspeed/inference/engine.py:39-636 (__init__ :44-187, _create_model_parallel_group :254-265,
_apply_injection_policy :408, CUDA graph :524-551, forward :584)."""
import os
import time

import torch
from torch.nn.modules import Module

from .. import comm as dist
from ..accelerator import get_accelerator
from ..module_inject.auto_tp import AutoTP
from ..module_inject.replace_module import replace_transformer_layer
from ..utils.logging import log_dist, logger
from .config import DeepSpeedInferenceConfig

INFERENCE_MODEL_TIMER = "model-forward-inference"


class InferenceEngine(Module):
    inference_mp_group = None
    inference_ep_group = None
    expert_mp_group = None

    def __init__(self, model, config: DeepSpeedInferenceConfig):
        super().__init__()
        self.module = model
        self._config = config
        self._get_model_config_generate(config)
        if hasattr(self.module, "generate"):
            self.generate = self._generate
        if hasattr(self.module, "config"):
            pass
        self.mp_world_size = config.tensor_parallel.tp_size
        self.checkpoint = config.checkpoint
        self.dtype = config.dtype
        self.injection_dict = config.injection_policy
        self.mp_group = config.tensor_parallel.tp_group
        self.mpu = config.tensor_parallel.mpu
        self.quantize_merge_count = 1
        self.quantization_scales = None
        self.ep_group = None
        self.expert_mp_group = None
        self.cuda_graph_created = False
        self.checkpoint_engine = None
        self._init_quantization_setting(None)
        self.model_profile_enabled = False
        self._model_times = []
        if config.enable_cuda_graph and config.tensor_parallel.tp_size > 1:
            pass  # graphs capture NCCL collectives fine on one NVSwitch box
        self._convert_to_dtype(config)
        if self.mpu:
            config.tensor_parallel.tp_size = dist.get_world_size(group=self.mpu.get_model_parallel_group())
            self.mp_group = self.mpu.get_model_parallel_group()
        elif config.tensor_parallel.tp_size > 1:
            self._create_model_parallel_group(config)
            config.tensor_parallel.tp_group = self.mp_group
        if isinstance(self.module, torch.nn.Module):
            moe, _ = False, 0
        if config.moe and getattr(config.moe, "enabled", False) and max(config.moe.moe_experts) > 1:
            self._create_ep_parallel_group(config.moe.moe_experts)
        if self.injection_dict:
            assert not config.replace_with_kernel_inject, "Cannot use both user specified injection policy and kernel injection"
            for client_module, injection_policy in self.injection_dict.items():
                assert issubclass(client_module, torch.nn.Module), f"{client_module} is not a subclass of torch.nn.Module"
                if isinstance(injection_policy, str):
                    config.injection_policy_tuple = (injection_policy, )
                else:
                    config.injection_policy_tuple = injection_policy
                layer_names = [name for name, _ in self.module.named_modules()]
                for policy in config.injection_policy_tuple:
                    if not any(name.endswith(policy) for name in layer_names):
                        raise ValueError(f"Injection policy layer'{policy}' not valid.")
                self._apply_injection_policy(config, client_module)
        else:
            self._kernel_injected = False
            if config.replace_with_kernel_inject:
                from .kernel_inject import KernelInjectedCausalLM, can_inject
                if can_inject(self.module):
                    # supported HF architecture: its weights are mapped onto the fused ragged transformer (blocked KV cache,
                    # fused norm/QKV/RoPE/attention/MLP kernels); tensor parallelism is handled inside that implementation
                    self.module = KernelInjectedCausalLM(self.module, dtype=config.dtype if config.dtype in (torch.bfloat16, torch.float16,
                                                                                                                torch.float32) else torch.bfloat16,
                                                         tp_size=config.tensor_parallel.tp_size, max_context=getattr(config, "max_out_tokens", None))
                    self._kernel_injected = True
                else:
                    # not a ragged-engine family (BERT, DistilBERT, GPT-Neo, BLOOM, CLIP, Megatron ...): convert layer by layer
                    from ..module_inject.replace_module import replace_with_kernel_policies
                    if config.dtype in (torch.float16, torch.bfloat16):
                        self.module.to(config.dtype)
                    n = replace_with_kernel_policies(self.module, config, getattr(self.module, "config", None))
                    if n:
                        self._kernel_injected = True
                        self._layer_injected = n
                        log_dist(f"kernel injection: {n} layers replaced by DeepSpeedTransformerInference blocks", ranks=[0])
                    else:
                        logger.warning("replace_with_kernel_inject: no layer policy matches this model; applying AutoTP")
            if config.tensor_parallel.tp_size > 1 and not self._kernel_injected:
                parser_dict = AutoTP.tp_parser(model)
                log_dist(f"AutoTP: {parser_dict}", ranks=[0])
                for client_module, injection_policy in parser_dict:
                    if isinstance(injection_policy, str):
                        config.injection_policy_tuple = (injection_policy, )
                    else:
                        config.injection_policy_tuple = injection_policy
                    self._apply_injection_policy(config, client_module)
        device = get_accelerator().current_device_name()
        is_meta = any(p.is_meta for p in self.module.parameters())
        if not is_meta:
            self.module.to(device)
        if config.tensor_parallel.tp_size > 1 and not is_meta and not getattr(self, "_kernel_injected", False) and not config.enable_cuda_graph:
            self._attach_fused_allreduce()
        if config.tensor_parallel.tp_size > 1:
            _rng_state = get_accelerator().get_rng_state().to(device)
            dist.broadcast(_rng_state, 0)
            get_accelerator().set_rng_state(_rng_state.cpu())
        self.local_cuda_graph = config.enable_cuda_graph and get_accelerator().device_name() == "cuda"
        self._graphs = {}

    def _attach_fused_allreduce(self):
        """Row-parallel layers of a single-node bf16 TP group run GEMM + all-reduce as one kernel over peer memory (one symmetric
        buffer and one FusedGemmAllReduce per distinct output width).  DSB200_FUSED_TP_AR=0 keeps NCCL; DSB200_FUSED_TP_MAX_TOKENS
        bounds the token count served by the fused path (larger inputs fall through to GEMM + NCCL).  Not used under CUDA-graph
        capture: the call epoch is a host-side counter."""
        import os
        from ..comm.comm import _single_node_job
        from ..module_inject.layers import LinearAllreduce
        if (os.environ.get("DSB200_FUSED_TP_AR", "1") == "0" or get_accelerator().device_name() != "cuda" or not _single_node_job()
                or self.mp_group is None or dist.get_world_size(self.mp_group) not in (2, 4, 8)):
            return
        layers = [m for m in self.module.modules() if isinstance(m, LinearAllreduce) and m.weight.dtype == torch.bfloat16
                  and m.weight.is_cuda and m.weight.shape[0] % 8 == 0]
        if not layers:
            return
        max_m = int(os.environ.get("DSB200_FUSED_TP_MAX_TOKENS", "4096"))
        try:
            from ..parallel.fused import FusedGemmAllReduce
            from ..parallel.symm import SymmetricBuffer
            self._fused_ar = {}
            for n in sorted({m.weight.shape[0] for m in layers}):
                sb = SymmetricBuffer(self.mp_group, FusedGemmAllReduce.bytes_needed(max_m, n))
                self._fused_ar[n] = FusedGemmAllReduce(sb, max_m, n)
                sb.barrier()
            for m in layers:
                m.attach_fused(self._fused_ar[m.weight.shape[0]])
            log_dist(f"tensor-parallel all-reduce fused into {len(layers)} row-parallel GEMMs (widths {sorted(self._fused_ar)})", ranks=[0])
        except Exception as e:  # no peer access between the ranks: NCCL
            logger.warning(f"fused GEMM + all-reduce unavailable ({e}); using NCCL")
            for m in layers:
                m._fused = None

    def profile_model_time(self, use_cuda_events=True):
        if not self.model_profile_enabled and not self._config.enable_cuda_graph:
            self.module.register_forward_pre_hook(self._pre_forward_hook)
            self.module.register_forward_hook(self._post_forward_hook)
        self.model_profile_enabled = True
        self.use_cuda_events = use_cuda_events and get_accelerator().device_name() == "cuda"
        if self.use_cuda_events:
            self.timers = __import__("deepspeedsyclsupport_b200.utils.timer", fromlist=["SynchronizedWallClockTimer"]).SynchronizedWallClockTimer()

    def _get_model_config_generate(self, config):
        self.config = getattr(self.module, "config", None) if config.config is None else config.config

    def _pre_forward_hook(self, module, *inputs, **kwargs):
        if self.use_cuda_events:
            self.timers(INFERENCE_MODEL_TIMER).start()
        else:
            get_accelerator().synchronize()
            self._start = time.time()

    def _post_forward_hook(self, module, input, output):
        if self.use_cuda_events:
            self.timers(INFERENCE_MODEL_TIMER).stop()
            elapsed_time = self.timers(INFERENCE_MODEL_TIMER).elapsed(reset=True)
        else:
            get_accelerator().synchronize()
            elapsed_time = (time.time() - self._start) * 1e3
        self._model_times.append(elapsed_time)

    def model_times(self):
        assert self.model_profile_enabled, "model profiling is not enabled"
        model_times = self._model_times
        if self._config.enable_cuda_graph and len(self._model_times) == 0:
            raise ValueError("Model times are empty and cuda graph is enabled. If this is a GPT-style model this combo is not supported.")
        self._model_times = []
        return model_times

    def _create_model_parallel_group(self, config):
        if InferenceEngine.inference_mp_group is None:
            if not dist.is_initialized():
                dist.init_distributed()
            local_rank = int(os.getenv("LOCAL_RANK", "0"))
            get_accelerator().set_device(local_rank)
            ranks = [i for i in range(config.tensor_parallel.tp_size)]
            self.mp_group = dist.new_group(ranks)
            InferenceEngine.inference_mp_group = self.mp_group
        else:
            self.mp_group = InferenceEngine.inference_mp_group

    def _create_ep_parallel_group(self, moe_experts):
        self.ep_group = {}
        self.expert_mp_group = {}
        moe_experts = moe_experts if isinstance(moe_experts, list) else [moe_experts]
        for e in moe_experts:
            self.ep_group.update({e: None})
            self.expert_mp_group.update({e: None})
        for moe_ep_size in self.ep_group.keys():
            num_ep_groups = dist.get_world_size() // moe_ep_size
            for i in range(num_ep_groups):
                ep_cnt = i * moe_ep_size
                size = dist.get_world_size() if moe_ep_size > dist.get_world_size() else moe_ep_size
                ranks = list(range(ep_cnt, ep_cnt + size))
                _ep_group = dist.new_group(ranks)
                if dist.get_rank() in ranks:
                    self.ep_group.update({moe_ep_size: _ep_group})

    def _init_quantization_setting(self, quantization_setting):
        self.quantize_bits = 8
        self.mlp_extra_grouping = False
        self.quantize_groups = 1
        if isinstance(quantization_setting, tuple):
            self.mlp_extra_grouping, self.quantize_groups = quantization_setting
        elif quantization_setting is not None:
            self.quantize_groups = quantization_setting

    def _apply_injection_policy(self, config, client_module=None):
        replace_transformer_layer(client_module, self.module, None, config, getattr(self.module, "config", None))

    def _convert_to_dtype(self, config):
        if not isinstance(self.module, torch.nn.Module):
            return
        if config.dtype == torch.half:
            self.module.half()
        elif config.dtype == torch.bfloat16:
            self.module.bfloat16()
        elif config.dtype == torch.float:
            self.module.float()

    # ---- CUDA graphs: one captured graph per input-shape signature (decode loops replay, no Python launch cost)
    def _signature(self, inputs, kwargs):
        sig = []
        for t in list(inputs) + [kwargs[k] for k in sorted(kwargs)]:
            if torch.is_tensor(t):
                sig.append((tuple(t.shape), t.dtype))
            else:
                sig.append(repr(t))
        return tuple(sig)

    def _create_cuda_graph(self, key, *inputs, **kwargs):
        cuda_stream = get_accelerator().Stream()
        cuda_stream.wait_stream(get_accelerator().current_stream())
        with get_accelerator().stream(cuda_stream):
            for i in range(3):
                ret = self.module(*inputs, **kwargs)
        get_accelerator().current_stream().wait_stream(cuda_stream)
        graph = get_accelerator().create_graph()
        static_inputs = inputs
        static_kwargs = kwargs
        with get_accelerator().capture_to_graph(graph):
            static_output = self.module(*static_inputs, **static_kwargs)
        self._graphs[key] = (graph, static_inputs, static_kwargs, static_output)
        self.cuda_graph_created = True

    def _graph_replay(self, key, *inputs, **kwargs):
        graph, static_inputs, static_kwargs, static_output = self._graphs[key]
        for i in range(len(inputs)):
            if torch.is_tensor(inputs[i]):
                static_inputs[i].copy_(inputs[i])
        for k in kwargs:
            if torch.is_tensor(kwargs[k]):
                static_kwargs[k].copy_(kwargs[k])
        get_accelerator().replay_graph(graph)
        return static_output

    def forward(self, *inputs, **kwargs):
        start = None
        if self.model_profile_enabled and get_accelerator().device_name() == "cuda" and self._config.enable_cuda_graph:
            get_accelerator().synchronize()
            start = time.time()
        if self.local_cuda_graph:
            key = self._signature(inputs, kwargs)
            if key in self._graphs:
                outputs = self._graph_replay(key, *inputs, **kwargs)
            else:
                self._create_cuda_graph(key, *inputs, **kwargs)
                outputs = self._graph_replay(key, *inputs, **kwargs)
        else:
            outputs = self.module(*inputs, **kwargs)
        if start is not None:
            get_accelerator().synchronize()
            self._model_times.append((time.time() - start) * 1e3)
        return outputs

    def _generate(self, *inputs, **kwargs):
        num_beams = 1
        if "generation_config" in kwargs:
            gen_config = kwargs["generation_config"]
            num_beams = getattr(gen_config, "num_beams", 1)
        if "num_beams" in kwargs:
            num_beams = kwargs["num_beams"]
        if num_beams > 1:
            raise NotImplementedError("DeepSpeed does not support `num_beams` > 1, if this is important to you please "
                                      "add your request to: https://github.com/microsoft/DeepSpeed/issues/2506")
        return self.module.generate(*inputs, **kwargs)

    # ------------------------------------------------------------------ ALiBi / attention-bias hooks for tensor parallelism
    # (reference: inference/engine.py:211-236 + module_inject/auto_tp_model_utils.py: HF BLOOM / MPT / Baichuan build an
    # alibi tensor for ALL heads; a tensor-parallel rank holds only its slice of the heads)
    def remove_mask_prepare_for_bloom(self):
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_prepare_attn_mask"):
            tr._prepare_attn_mask = lambda attention_mask, *args, **kwargs: attention_mask

    def _tp_rank_world(self):
        if self.mp_group is not None and dist.is_initialized():
            return dist.get_rank(group=self.mp_group), dist.get_world_size(group=self.mp_group)
        return 0, 1

    def build_alibi_tensor(self):
        from ..module_inject.auto_tp_model_utils import build_bloom_alibi_tensor, build_mpt_alibi_tensor, get_alibi_mask
        tr = getattr(self.module, "transformer", None)
        if tr is not None:
            if hasattr(tr, "build_alibi_tensor"):
                tr.build_alibi_tensor = build_bloom_alibi_tensor
            if hasattr(tr, "build_mpt_alibi_tensor"):
                tr.build_mpt_alibi_tensor_orig = tr.build_mpt_alibi_tensor
                tr.__class__.build_mpt_alibi_tensor = build_mpt_alibi_tensor
        inner = getattr(self.module, "model", None)
        if inner is not None and hasattr(inner, "get_alibi_mask"):
            inner.get_alibi_mask_orig = inner.get_alibi_mask
            inner.__class__.get_alibi_mask = get_alibi_mask

    def build_attn_bias(self):
        from ..module_inject.auto_tp_model_utils import build_mpt_atten_bias_tensor
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_attn_bias"):
            tr._attn_bias_orig = tr._attn_bias
            tr.__class__._attn_bias = build_mpt_atten_bias_tensor

    # ------------------------------------------------------------------ checkpoints
    def load_model_with_checkpoint(self, r_module):
        """Fill `r_module` (possibly built on the meta device) from `config.checkpoint`: a DeepSpeed checkpoint description
        ({"type", "checkpoints": [...], "version"} dict or json path), a directory of torch state-dict shards, or one file.
        Tensor-parallel merge / split of Megatron-style shards goes through SDLoaderFactory."""
        import glob
        import os
        from ..runtime.checkpoint_engine.torch_checkpoint_engine import TorchCheckpointEngine
        from ..runtime.state_dict_factory import SDLoaderFactory
        self.checkpoint_engine = self.checkpoint_engine or TorchCheckpointEngine()
        ckpt = self.checkpoint
        rank, world = self._tp_rank_world()
        if isinstance(ckpt, dict) or (isinstance(ckpt, str) and ckpt.endswith(".json")):
            loader = SDLoaderFactory.get_sd_loader_json(ckpt, self.checkpoint_engine)
            if isinstance(loader, tuple) or isinstance(loader, list):  # (type, list, version) for non-Megatron descriptions
                files = loader[1] if len(loader) > 1 else []
                sds = [self.checkpoint_engine.load(f, map_location="cpu") for f in files]
            else:
                _, sd, _ = loader.load(world, rank)
                sds = [loader.get_module(sd) if hasattr(loader, "get_module") else sd]
        else:
            files = [ckpt] if os.path.isfile(ckpt) else sorted(glob.glob(os.path.join(ckpt, "*.pt")) + glob.glob(os.path.join(ckpt, "*.bin")))
            assert files, f"no checkpoint files under {ckpt}"
            sds = [self.checkpoint_engine.load(f, map_location="cpu") for f in files]
        merged = {}
        for sd in sds:
            merged.update(sd.get("module", sd.get("model", sd)) if isinstance(sd, dict) else sd)
        if any(p.is_meta for p in r_module.parameters()):
            r_module.to_empty(device=get_accelerator().current_device_name())
        missing, unexpected = r_module.load_state_dict(merged, strict=False)
        log_dist(f"load_model_with_checkpoint: {len(merged)} tensors, {len(missing)} missing, {len(unexpected)} unexpected", ranks=[0])
        return r_module

    def destroy(self):
        """Release graphs / workspaces held by the engine (reference: InferenceEngine.destroy frees the kernel workspace)."""
        self._graphs = {}
        self.cuda_graph_created = False
        inner = getattr(self.module, "engine", None)  # kernel-injected model: flush every tracked sequence
        if inner is not None and hasattr(inner, "_state_manager"):
            for uid in list(getattr(inner._state_manager, "tracked_sequences", {}).keys()):
                inner.flush(uid)
