"""offload_param / offload_optimizer config (parity: deepspeed/runtime/zero/offload_config.py:12-97)."""
from enum import Enum
from pathlib import Path
from typing import Optional

from pydantic import Field, model_validator

from ..config_utils import DeepSpeedConfigModel, pp_int


class OffloadDeviceEnum(str, Enum):
    none = "none"
    cpu = "cpu"
    nvme = "nvme"


class DeepSpeedZeroOffloadParamConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = "none"
    nvme_path: Optional[Path] = None
    buffer_count: int = Field(5, ge=0)
    buffer_size: int = Field(pp_int(1e8), ge=0)
    max_in_cpu: int = Field(pp_int(1e9), ge=0)
    pin_memory: bool = False


class DeepSpeedZeroOffloadOptimizerConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = "none"
    nvme_path: Optional[Path] = None
    buffer_count: int = Field(4, ge=0)
    pin_memory: bool = False
    pipeline_read: bool = False
    pipeline_write: bool = False
    fast_init: bool = False
    ratio: float = Field(1.0, ge=0.0, le=1.0)  # Offload++ / Twin-Flow: fraction of sub-groups stepped on the host
    pipeline: bool = False

    @model_validator(mode="after")
    def set_pipeline(self):
        object.__setattr__(self, "pipeline", bool(self.pipeline_read or self.pipeline_write))
        return self
