"""Containers: build a DeepSpeedTransformerInference block from the tensors a policy exposes (parity:
deepspeed/module_inject/containers/base.py:18-320).

Flow (driven by module_inject/replace_module.replace_transformer_layer):
    c = DS_XContainer(policy=X_Policy(layer), config=inference_config, model_config=hf_config, layer_id=i)
    c.create_module()               # block with empty parameters of this rank's (tensor-parallel) shapes
    c.initialize_tensors()          # read the client's tensors through the policy
    c.transpose()                   # Conv1D layouts -> [out, in]
    c.apply_tensor_parallelism(mp_replace)   # copy (slices of) the tensors into the block
    c.copy_data_to_new_module()     # norms (never sliced)
The block's kernels are this repo's (tcgen05 GEMM, fused norms); `dtype` int8 is served as 16-bit weights."""
from abc import ABC

import torch

from ...model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
from ...ops.transformer.inference.config import DeepSpeedInferenceConfig
from ...utils.types import GATED_ACTIVATION_TYPES, ActivationFuncType
from ..policy import DEFAULT_INTERMEDIATE_SIZE, transpose


class BaseConvolutionContainer(ABC):
    """Marker base of the diffusers (UNet / VAE) containers, which wrap whole modules rather than transformer layers."""


class BaseTransformerContainer(ABC):

    def __init__(self, policy, config, model_config, layer_id, child):
        self.policy = policy
        self.config = config
        self.model_config = model_config
        self.layer_id = layer_id
        self.child = child
        self.megatron_v2 = self.policy.is_megatron_v2
        self.scale_attention = self.policy.scale_attention
        self.ckpt_load_enabled = False
        self._h = None
        self.hidden_size = self.num_attention_heads = None
        self.mp_size = self.config.tensor_parallel.tp_size
        self.pre_layer_norm = self.model_config.do_layer_norm_before if hasattr(self.model_config, "do_layer_norm_before") else \
            self.policy.pre_attn_norm
        self.dtype = self.config.dtype
        self.intermediate_size = None
        self.fp16 = self.dtype in (torch.float16, torch.int8)
        self.attn_linear_layer = self.policy.linear_layer
        self.mlp_linear_layer = self.policy.linear_layer
        self.return_tuple = self.config.return_tuple
        self.triangular_masking = True
        self.local_attention = ((self.model_config.attention_layers[self.layer_id] == "local") if hasattr(self.model_config, "attention_layers")
                                else False)
        self.window_size = getattr(self.model_config, "window_size", 1)
        self.mlp_act_func_type = self.policy.mlp_act_func_type
        self.norm_type = self.policy.norm_type
        self.training_mp_size = self.config.training_mp_size
        self.bigscience_bloom = False
        self.max_out_tokens = self.config.max_out_tokens
        self.min_out_tokens = self.config.min_out_tokens
        self.scale_attn_by_inverse_layer_idx = getattr(self.config, "scale_attn_by_inverse_layer_idx", False)
        self.use_mup = self.policy.use_mup
        self.return_single_tuple = False
        self.rotary_dim = self.model_config.rotary_dim if hasattr(self.model_config, "rotary_dim") else \
            (self.child.attention.rotary_ndims if hasattr(self.child, "attention") and hasattr(self.child.attention, "rotary_ndims") else -1)
        self.mlp_after_attn = (self.rotary_dim is None or self.rotary_dim < 0)
        self.rotate_half = self.rotate_every_two = False
        self.rope_theta = getattr(self.model_config, "rope_theta", 10000)
        self.num_kv = -1
        # tensors read from the client layer
        self.qkvw = self.qkvb = self.dense_w = self.dense_b = None
        self._h4h_w = self._h4h_b = self._4hh_w = self._4hh_b = None
        self.attn_nw = self.attn_nb = self.input_nw = self.input_nb = None
        self.mp_group = None
        self.use_triton = False
        self.module = None
        self.ds_model_config = None

    # ---- configuration
    def create_ds_model_config(self):
        self.set_hidden_heads(*self.policy.get_hidden_heads())
        assert self.num_attention_heads % self.mp_size == 0, \
            "To run the model parallel across the GPUs, the attention_heads require to be divisible by the world_size!" \
            "This is because the attention computation is partitioned evenly among the parallel GPUs."
        self.ds_model_config = DeepSpeedInferenceConfig(
            hidden_size=self.hidden_size, intermediate_size=self.intermediate_size, heads=self.num_attention_heads,
            layer_norm_eps=self.layernorm_epsilon, dtype=torch.half if self.dtype == torch.int8 else self.dtype, pre_layer_norm=self.pre_layer_norm,
            norm_type=self.norm_type, mp_size=self.mp_size, return_tuple=self.return_tuple, triangular_masking=self.triangular_masking,
            local_attention=self.local_attention, window_size=self.window_size, rotary_dim=self.rotary_dim if self.rotary_dim else -1,
            rotate_half=self.rotate_half, rotate_every_two=self.rotate_every_two, mlp_after_attn=self.mlp_after_attn,
            mlp_act_func_type=self.mlp_act_func_type, training_mp_size=self.training_mp_size, bigscience_bloom=self.bigscience_bloom,
            max_out_tokens=self.max_out_tokens, min_out_tokens=self.min_out_tokens,
            scale_attn_by_inverse_layer_idx=self.scale_attn_by_inverse_layer_idx, use_mup=self.use_mup, return_single_tuple=self.return_single_tuple,
            scale_attention=self.scale_attention, num_kv=self.num_kv, rope_theta=self.rope_theta)
        return self.ds_model_config

    def set_hidden_heads(self, hidden_size, num_attention_heads, epsilon, intermediate_size):
        self.hidden_size = hidden_size
        self.num_attention_heads = num_attention_heads
        self.layernorm_epsilon = epsilon
        self.intermediate_size = intermediate_size if intermediate_size != DEFAULT_INTERMEDIATE_SIZE else 4 * hidden_size

    def create_module(self, config=None):
        """The block for this architecture (subclasses choose the DeepSpeed*Inference class)."""
        cfg = config if config is not None else (self.ds_model_config or self.create_ds_model_config())
        self.module = DeepSpeedTransformerInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    # ---- reading the client layer
    def initialize_tensors(self, enable_training=False):
        self.set_attention(*self.policy.attention())
        self.set_mlp(*self.policy.mlp())
        self.set_layernorm(*self.policy.layernorm())

    def set_attention(self, qkvw, qkvb, dense_w, dense_b):
        self.qkvw, self.qkvb, self.dense_w, self.dense_b = qkvw, qkvb, dense_w, dense_b

    def set_mlp(self, _h4h_w, _h4h_b, _4hh_w, _4hh_b):
        self._h4h_w, self._h4h_b, self._4hh_w, self._4hh_b = _h4h_w, _h4h_b, _4hh_w, _4hh_b

    def set_layernorm(self, attn_nw, attn_nb, input_nw, input_nb):
        self.attn_nw, self.attn_nb, self.input_nw, self.input_nb = attn_nw, attn_nb, input_nw, input_nb

    def set_dtype(self, fp16=False):
        self.fp16 = fp16

    def set_tensor_parallel_config(self, mp_size, mp_group):
        self.mp_size, self.mp_group = mp_size, mp_group

    def set_quantization_config(self, quantizer):
        self.quantizer = quantizer

    # ---- layout
    def transpose(self):
        """Conv1D-style layers store [in, out]: bring every projection to [out, in]."""
        if not self.attn_linear_layer:
            self.qkvw, self.dense_w = transpose(self.qkvw.data), transpose(self.dense_w.data)
        if not self.mlp_linear_layer:
            self._h4h_w, self._4hh_w = transpose(self._h4h_w.data), transpose(self._4hh_w.data)

    # ---- filling the block
    def apply_tensor_parallelism(self, mp_replace, **kwargs):
        self.attention_qkv_mp(mp_replace)
        self.attention_o_mp(mp_replace)
        self.mlp_inter_mp(mp_replace)
        self.mlp_output_mp(mp_replace)

    def _optional(self, owner, name, src, mp_replace, fused=0):
        """Copy `src` into owner.<name>; a bias the client does not have becomes zeros."""
        dst = getattr(owner, name)
        if src is None:
            with torch.no_grad():
                dst.zero_()
            return
        setattr(owner, name, mp_replace.strided_copy(dst, src, num_splits=fused) if fused else mp_replace.copy(dst, src))

    def attention_qkv_mp(self, mp_replace, reversed_dim=False):
        a = self.module.attention
        self._optional(a, "attn_qkvw", self.qkvw, mp_replace, fused=3)
        self._optional(a, "attn_qkvb", self.qkvb, mp_replace, fused=3)

    def attention_o_mp(self, mp_replace, reversed_dim=False):
        a = self.module.attention
        self._optional(a, "attn_ow", self.dense_w, mp_replace)
        self._optional(a, "attn_ob", self.dense_b, mp_replace)

    def mlp_inter_mp(self, mp_replace, reversed_dim=False):
        m = self.module.mlp
        fused = 2 if self.mlp_act_func_type in GATED_ACTIVATION_TYPES else 0
        self._optional(m, "inter_w", self._h4h_w, mp_replace, fused=fused)
        self._optional(m, "inter_b", self._h4h_b, mp_replace, fused=fused)

    def mlp_output_mp(self, mp_replace, reversed_dim=False):
        m = self.module.mlp
        self._optional(m, "output_w", self._4hh_w, mp_replace)
        self._optional(m, "output_b", self._4hh_b, mp_replace)

    def copy_data_to_new_module(self):
        """Norm parameters are replicated: straight copies.  A norm the architecture lacks keeps gamma = 1, beta = 0."""

        def put(dst, src, fill):
            with torch.no_grad():
                if src is None:
                    dst.fill_(fill)
                else:
                    dst.copy_(src.to(dst.dtype))

        put(self.module.mlp.attn_nw, self.attn_nw, 1.0)
        put(self.module.mlp.attn_nb, self.attn_nb, 0.0)
        put(self.module.norm_w, self.input_nw, 1.0)
        put(self.module.norm_b, self.input_nb, 0.0)

    def get_all_params(self):
        return [p for p in (self.qkvw, self.qkvb, self.dense_w, self.dense_b, self._h4h_w, self._h4h_b, self._4hh_w, self._4hh_b,
                            self.attn_nw, self.attn_nb, self.input_nw, self.input_nb) if p is not None]

    def align_merged_qkv(self):
        pass

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        """Fill `module` straight from a checkpoint state dict (meta-tensor loading); architecture containers that support it
        mix in features.meta_tensor.MetaTensorContainer and implement this."""
        raise NotImplementedError(f"{type(self).__name__} does not load from checkpoints directly")
