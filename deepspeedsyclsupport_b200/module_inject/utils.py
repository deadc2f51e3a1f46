"""(parity: deepspeed/module_inject/utils.py)"""
from .policy import DSPolicy


def policy_to_ds_container(**kwargs):
    """The container class that goes with a policy instance, instantiated with the given arguments."""
    from .containers import (BLOOMLayerPolicy, DS_BERTContainer, DS_BloomContainer, DS_CLIPContainer, DS_DistilBERTContainer, DS_GPT2Container,
                             DS_GPTJContainer, DS_GPTNEOContainer, DS_GPTNEOXContainer, DS_InternLMContainer, DS_LLAMA2Container,
                             DS_LLAMAContainer, DS_MegatronGPTContainer, DS_MegatronGPTMoEContainer, DS_OPTContainer, GPTNEOXLayerPolicy,
                             HFBertLayerPolicy, HFCLIPLayerPolicy, HFDistilBertLayerPolicy, HFGPT2LayerPolicy, HFGPTJLayerPolicy,
                             HFGPTNEOLayerPolicy, HFOPTLayerPolicy, InternLMLayerPolicy, LLAMA2LayerPolicy, LLAMALayerPolicy,
                             MegatronLayerPolicy, MegatronMoELayerPolicy)
    policy_to_container = {
        HFGPT2LayerPolicy: DS_GPT2Container, HFBertLayerPolicy: DS_BERTContainer, BLOOMLayerPolicy: DS_BloomContainer,
        HFGPTJLayerPolicy: DS_GPTJContainer, HFGPTNEOLayerPolicy: DS_GPTNEOContainer, GPTNEOXLayerPolicy: DS_GPTNEOXContainer,
        HFOPTLayerPolicy: DS_OPTContainer, MegatronLayerPolicy: DS_MegatronGPTContainer, MegatronMoELayerPolicy: DS_MegatronGPTMoEContainer,
        HFDistilBertLayerPolicy: DS_DistilBERTContainer, HFCLIPLayerPolicy: DS_CLIPContainer, LLAMALayerPolicy: DS_LLAMAContainer,
        LLAMA2LayerPolicy: DS_LLAMA2Container, InternLMLayerPolicy: DS_InternLMContainer,
    }
    policy = kwargs.get("policy")
    assert policy is not None, "Policy cannot be None"
    policy_type = type(policy)
    if policy_type not in policy_to_container:
        raise ValueError(f"Policy type {policy_type} not supported")
    return policy_to_container[policy_type](**kwargs)
