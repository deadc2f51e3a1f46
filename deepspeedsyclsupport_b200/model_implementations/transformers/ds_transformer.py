"""`DeepSpeedTransformerInference`: one transformer block for inference with the reference's module / parameter names
(model_implementations/transformers/ds_transformer.py:23-199): `norm_w / norm_b` (input norm), `attention` (DeepSpeedSelfAttention
with its KV cache), `mlp` (DeepSpeedMLP with the post-attention norm).  Runs on this framework's kernels: tcgen05 GEMM for the
projections, fused norm / bias-activation kernels; attention over the cached K / V.
The per-architecture subclasses of the reference (ds_bert / ds_gpt / ds_opt / ds_llama2 / ds_megatron_gpt / ds_bloom) live in
their own modules next to this one; they differ only in what the containers configure."""
import torch
from torch import nn

from ...accelerator import get_accelerator
from ...ops.transformer.inference import DeepSpeedMLP, DeepSpeedSelfAttention
from ...ops.transformer.inference.ds_mlp import block_norm
from ...utils.logging import log_dist


class DeepSpeedTransformerInference(nn.Module):
    layer_id = 0

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerInference.layer_id
        DeepSpeedTransformerInference.layer_id += 1
        if DeepSpeedTransformerInference.layer_id == 1:
            log_dist(f"DeepSpeed-Inference config: {self.config.__dict__}", [0])
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        self.attention = DeepSpeedSelfAttention(config, mp_group, quantize_scales, quantize_groups, merge_count)
        self.mlp = DeepSpeedMLP(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
        dev = get_accelerator().current_device_name()
        if config.set_empty_params:
            self.norm_w = self.norm_b = None
        else:
            self.norm_w = nn.Parameter(torch.empty(config.hidden_size, dtype=dt, device=dev), requires_grad=False)
            self.norm_b = nn.Parameter(torch.empty(config.hidden_size, dtype=dt, device=dev), requires_grad=False)
        self.layer_past = None
        # parallel-residual blocks whose MLP has its own LayerNorm of the block input (GPT-NeoX); GPT-J style blocks feed the
        # MLP the same normalised input as the attention
        self.parallel_mlp_norm = False
        # client layers that return the bare hidden states even when caching (newer HF decoder layers): presents stay internal
        self.tensor_output_only = False

    @classmethod
    def reset_cache(cls):
        """Start a new generation: every layer drops its KV cache on its next multi-token call (the reference resets a shared
        workspace); kept as a class method for API parity."""
        cls._reset_epoch = getattr(cls, "_reset_epoch", 0) + 1

    def forward(self, input=None, input_mask=None, attention_mask=None, attn_mask=None, head_mask=None, layer_past=None, get_key_value=False,
                get_present=False, encoder_output=None, enc_dec_attn_mask=None, x=None, encoder_hidden_states=None, encoder_attention_mask=None,
                use_cache=False, alibi=None, output_attentions=False, layer_head_mask=None, past_key_value=None, **kwargs):
        if x is not None:
            input = x
        if "hidden_states" in kwargs:
            input = kwargs["hidden_states"]
        mask = attention_mask if attention_mask is not None else (attn_mask if attn_mask is not None else input_mask)
        get_present = get_present or get_key_value or use_cache
        if isinstance(input, tuple):
            mask = input[1] if mask is None else mask
            input = input[0]
        in_dtype = input.dtype
        cfg = self.config
        if cfg.dtype in (torch.float16, torch.bfloat16, torch.int8) and input.dtype == torch.float32:
            input = input.to(torch.half if cfg.dtype == torch.int8 else cfg.dtype)
        epoch = getattr(DeepSpeedTransformerInference, "_reset_epoch", 0)
        if getattr(self, "_seen_epoch", 0) != epoch:
            self.attention.reset_cache()
            self._seen_epoch = epoch
        past = layer_past if layer_past is not None else past_key_value
        if past is not None and not isinstance(past, (tuple, list)):
            past = None  # a framework-managed cache object (HF Cache): the block keeps its own K / V, as the reference's does
        with torch.no_grad():
            hn = block_norm(input, self.norm_w, self.norm_b, cfg) if cfg.pre_layer_norm else input
            attn_out, key, value = self.attention(hn, mask, head_mask, past, get_present, encoder_hidden_states, encoder_attention_mask,
                                                  output_attentions, self.norm_w, self.norm_b, alibi)
            if not cfg.mlp_after_attn:
                # parallel block (GPT-J / NeoX style): attention and MLP read the same normed input
                mlp_in = block_norm(input, self.mlp.attn_nw, self.mlp.attn_nb, cfg) if self.parallel_mlp_norm else hn
                output = input + attn_out + self.mlp.mlp_only(mlp_in)
            elif cfg.pre_layer_norm:
                output = self.mlp(attn_out, input)
            else:  # post-LN (BERT): LN_attn(x + attn) -> + mlp -> LN_out
                output = block_norm(self.mlp(attn_out, input), self.norm_w, self.norm_b, cfg)
        output = output.to(in_dtype)
        if self.tensor_output_only:
            return output
        presents = (key, value)
        if get_present:
            output = (output, presents)
        if cfg.return_single_tuple:
            return (output, )
        if cfg.return_tuple:
            return output if isinstance(output, tuple) else (output, mask)
        return output
