"""DeepSpeedInferenceConfig (parity: deepspeed/inference/config.py:128-304 -- same field names and aliases)."""
from enum import Enum
from typing import Any, Dict, Optional, Union

import torch
from pydantic import Field, field_validator

from ..runtime.config_utils import DeepSpeedConfigModel


class DtypeEnum(Enum):
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    @classmethod
    def from_any(cls, v):
        for m in cls:
            if v in m.value or v is m:
                return m
        raise ValueError(f"unknown dtype {v}")


class MoETypeEnum(str, Enum):
    residual = "residual"
    standard = "standard"


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    enabled: bool = True
    tp_size: int = 1
    mpu: Any = None
    tp_group: Any = None


class DeepSpeedMoEConfig(DeepSpeedConfigModel):
    enabled: bool = True
    ep_size: int = 1
    moe_experts: list = Field([1], alias="num_experts")
    type: MoETypeEnum = MoETypeEnum.standard
    ep_mp_group: Any = None
    ep_group: Any = Field(None, alias="expert_group")


class QuantTypeEnum(str, Enum):
    asym = "asymmetric"
    sym = "symmetric"


class BaseQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True
    num_bits: int = 8
    q_type: QuantTypeEnum = QuantTypeEnum.sym
    q_groups: int = 1


class WeightQuantConfig(BaseQuantConfig):
    enabled: bool = True
    quantized_initialization: Dict = {}
    post_init_quant: Dict = {}


class ActivationQuantConfig(BaseQuantConfig):
    enabled: bool = True


class QKVQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True


class QuantizationConfig(DeepSpeedConfigModel):
    enabled: bool = True
    activation: ActivationQuantConfig = ActivationQuantConfig()
    weight: WeightQuantConfig = WeightQuantConfig()
    qkv: QKVQuantConfig = QKVQuantConfig()


class InferenceCheckpointConfig(DeepSpeedConfigModel):
    checkpoint_dir: Optional[str] = None
    save_mp_checkpoint_path: Optional[str] = None
    base_dir: Optional[str] = None


class DeepSpeedInferenceConfig(DeepSpeedConfigModel):
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")
    dtype: Any = torch.float16
    tensor_parallel: DeepSpeedTPConfig = Field(DeepSpeedTPConfig(), alias="tp")
    enable_cuda_graph: bool = False
    use_triton: bool = False
    triton_autotune: bool = False
    zero: Dict = {}
    triangular_masking: bool = Field(True, alias="tm")
    moe: Union[bool, DeepSpeedMoEConfig] = {}
    quant: QuantizationConfig = {}
    checkpoint: Optional[Union[str, Dict]] = None
    base_dir: str = ""
    set_empty_params: bool = False
    save_mp_checkpoint_path: Optional[str] = None
    checkpoint_config: InferenceCheckpointConfig = Field({}, alias="ckpt_config")
    return_tuple: bool = True
    training_mp_size: int = 1
    replace_method: str = Field("auto", json_schema_extra={"deprecated": True, "deprecated_msg": "This parameter is no longer needed, please remove from your call to DeepSpeed-inference"})
    injection_policy: Optional[Dict] = Field(None, alias="injection_dict")
    injection_policy_tuple: Optional[tuple] = None
    config: Optional[Dict] = Field(None, alias="args")
    max_out_tokens: int = Field(1024, alias="max_tokens")
    min_out_tokens: int = Field(1, alias="min_tokens")
    transposed_mode: bool = Field(False, alias="transposed_mode")
    mp_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.tp_size"})
    mpu: Any = Field(None, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.mpu"})
    ep_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "moe.ep_size"})
    ep_group: Any = Field(None, alias="expert_group", json_schema_extra={"deprecated": True, "new_param": "moe.ep_group"})
    ep_mp_group: Any = Field(None, alias="expert_mp_group", json_schema_extra={"deprecated": True, "new_param": "moe.ep_mp_group"})
    moe_experts: list = Field([1], json_schema_extra={"deprecated": True, "new_param": "moe.moe_experts"})
    moe_type: MoETypeEnum = Field(MoETypeEnum.standard, json_schema_extra={"deprecated": True, "new_param": "moe.type"})

    @field_validator("dtype", mode="before")
    @classmethod
    def _dtype(cls, v):
        if isinstance(v, torch.dtype):
            return v
        return DtypeEnum.from_any(v).value[0]

    @field_validator("moe", mode="before")
    @classmethod
    def moe_backward_compat(cls, field_value):
        if isinstance(field_value, bool):
            return DeepSpeedMoEConfig(moe=field_value) if False else DeepSpeedMoEConfig(enabled=field_value)
        if isinstance(field_value, dict):
            return DeepSpeedMoEConfig(**field_value)
        return field_value

    @field_validator("quant", mode="before")
    @classmethod
    def _quant(cls, v):
        return QuantizationConfig(**v) if isinstance(v, dict) else v

    @field_validator("checkpoint_config", mode="before")
    @classmethod
    def _ckpt(cls, v):
        return InferenceCheckpointConfig(**v) if isinstance(v, dict) else v
