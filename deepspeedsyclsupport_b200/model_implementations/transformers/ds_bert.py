"""(parity: deepspeed/model_implementations/transformers/ds_bert.py)"""
from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedBERTInference(DeepSpeedTransformerInference):
    """BERT encoder block: bidirectional (no triangular mask), post-LN, returns a tuple."""

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
