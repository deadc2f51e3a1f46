"""One place that decides which implementation serves each building block of a model (parity:
inference/v2/modules/heuristics.py:36-179).  Decisions here depend on the engine configuration, not on the model: the
quantisation mode selects the fp8 tensor-core linear; everything else currently has one B200 implementation."""
from ..config_v2 import RaggedInferenceEngineConfig
from ..inference_utils import NormTypeEnum
from . import implementations  # noqa: F401  (registers the implementations)
from .configs import DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig, DSUnembedConfig
from .interfaces import (DSEmbeddingBase, DSEmbeddingRegistry, DSLinearBase, DSLinearRegistry, DSMoEBase, DSMoERegistry, DSPostNormBase,
                         DSPostNormRegistry, DSPreNormBase, DSPreNormRegistry, DSSelfAttentionBase, DSSelfAttentionRegistry, DSUnembedBase,
                         DSUnembedRegistry)
from .module_registry import ConfigBundle


def _quant_mode(engine_config) -> str:
    q = getattr(engine_config, "quantization", None) if engine_config is not None else None
    return getattr(q, "quantization_mode", "") or ""


def instantiate_attention(attention_config: DSSelfAttentionConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSSelfAttentionBase:
    return DSSelfAttentionRegistry.instantiate_config(ConfigBundle(name="dense_blocked_attention", config=attention_config))


def instantiate_embed(embed_config: DSEmbeddingsConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSEmbeddingBase:
    return DSEmbeddingRegistry.instantiate_config(ConfigBundle(name="ragged_embedding", config=embed_config))


def instantiate_linear(linear_config: DSLinearConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSLinearBase:
    name = "blas_fp_linear"
    mode = _quant_mode(engine_config)
    if mode == "wf8af8" and DSLinearRegistry.registry["fp8_linear"].supports_config(linear_config):
        name = "fp8_linear"
    elif mode == "wf8af8_block" and DSLinearRegistry.registry["fp8_block_linear"].supports_config(linear_config):
        name = "fp8_block_linear"
    elif mode == "wf8af8_mx" and DSLinearRegistry.registry["mxfp8_linear"].supports_config(linear_config):
        name = "mxfp8_linear"
    return DSLinearRegistry.instantiate_config(ConfigBundle(name=name, config=linear_config))


def instantiate_moe(moe_config: DSMoEConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSMoEBase:
    bundle = ConfigBundle(name="cutlass_multi_gemm_moe", config=moe_config, implementation_config={"weight_dtype": moe_config.input_dtype})
    return DSMoERegistry.instantiate_config(bundle)


def instantiate_post_norm(norm_config: DSNormConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSPostNormBase:
    return DSPostNormRegistry.instantiate_config(ConfigBundle(name="cuda_post_ln", config=norm_config))


def instantiate_pre_norm(norm_config: DSNormConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSPreNormBase:
    kind = NormTypeEnum(norm_config.type)
    name = "cuda_pre_ln" if kind == NormTypeEnum.LayerNorm else "cuda_pre_rms"
    return DSPreNormRegistry.instantiate_config(ConfigBundle(name=name, config=norm_config))


def instantiate_unembed(unembed_config: DSUnembedConfig, engine_config: RaggedInferenceEngineConfig = None) -> DSUnembedBase:
    return DSUnembedRegistry.instantiate_config(ConfigBundle(name="ragged_unembed", config=unembed_config))
