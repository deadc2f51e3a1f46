from .bert import DS_BERTContainer, HFBertLayerPolicy
from .bloom import BLOOMLayerPolicy, DS_BloomContainer
from .clip import DS_CLIPContainer, HFCLIPLayerPolicy
from .distil_bert import DS_DistilBERTContainer, HFDistilBertLayerPolicy
from .gpt2 import DS_GPT2Container, HFGPT2LayerPolicy
from .gptj import DS_GPTJContainer, HFGPTJLayerPolicy
from .gptneo import DS_GPTNEOContainer, HFGPTNEOLayerPolicy
from .gptneox import DS_GPTNEOXContainer, GPTNEOXLayerPolicy
from .internlm import DS_InternLMContainer, InternLMLayerPolicy
from .llama import DS_LLAMAContainer, LLAMALayerPolicy
from .llama2 import DS_LLAMA2Container, LLAMA2LayerPolicy
from .megatron_gpt import DS_MegatronGPTContainer, MegatronLayerPolicy
from .megatron_gpt_moe import DS_MegatronGPTMoEContainer, MegatronMoELayerPolicy
from .opt import DS_OPTContainer, HFOPTLayerPolicy
from .unet import UNetPolicy
from .vae import VAEPolicy
