"""RaggedInferenceEngineConfig (parity: inference/v2/config_v2.py)."""
from pydantic import Field

from ...runtime.config_utils import DeepSpeedConfigModel
from .ragged.manager_configs import DSStateManagerConfig


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    tp_size: int = 1


class QuantizationConfig(DeepSpeedConfigModel):
    # "" (16-bit weights) | "wf8af8" (e4m3 weights per output feature x e4m3 activations per token)
    # | "wf8af8_block" (block-scaled: one fp32 scale per 128 input features on both operands, the ZeRO++ wire format)
    # | "wf8af8_mx" (MXFP8: one UE8M0 scale per 32 input features, applied by the tensor core: tcgen05 mxf8f6f4.block_scale)
    quantization_mode: str = ""


class RaggedInferenceEngineConfig(DeepSpeedConfigModel):
    tensor_parallel: DeepSpeedTPConfig = Field(DeepSpeedTPConfig(), alias="tp")
    state_manager: DSStateManagerConfig = Field(DSStateManagerConfig(), alias="manager")
    quantization: QuantizationConfig = QuantizationConfig()
