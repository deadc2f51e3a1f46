"""(parity: deepspeed/model_implementations/transformers/ds_llama2.py)"""
from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedLlama2Inference(DeepSpeedTransformerInference):
    """Llama / Llama-2 block: RMSNorm, gated SiLU MLP, rotate-half rotary, GQA through config.num_kv."""

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
