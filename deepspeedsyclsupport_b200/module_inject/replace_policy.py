"""The layer policies tried, in order, when kernel injection looks for convertible layers (parity:
deepspeed/module_inject/replace_policy.py)."""
from .containers import (BLOOMLayerPolicy, GPTNEOXLayerPolicy, HFBertLayerPolicy, HFCLIPLayerPolicy, HFDistilBertLayerPolicy, HFGPT2LayerPolicy,
                         HFGPTJLayerPolicy, HFGPTNEOLayerPolicy, HFOPTLayerPolicy, InternLMLayerPolicy, LLAMA2LayerPolicy, LLAMALayerPolicy,
                         MegatronLayerPolicy, UNetPolicy, VAEPolicy)

# transformer-based policies
replace_policies = [HFBertLayerPolicy, HFGPTNEOLayerPolicy, GPTNEOXLayerPolicy, HFGPTJLayerPolicy, MegatronLayerPolicy, HFGPT2LayerPolicy,
                    BLOOMLayerPolicy, HFOPTLayerPolicy, HFCLIPLayerPolicy, HFDistilBertLayerPolicy, LLAMALayerPolicy, LLAMA2LayerPolicy,
                    InternLMLayerPolicy]

# non-transformer-based policies
generic_policies = [UNetPolicy, VAEPolicy]
