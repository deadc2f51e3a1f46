"""comms_logger config block (parity: deepspeed/comm/config.py:18-33)."""
from typing import List

from ..runtime.config_utils import DeepSpeedConfigModel

COMMS_LOGGER = "comms_logger"


class CommsLoggerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    prof_all: bool = True
    prof_ops: List[str] = []
    verbose: bool = False
    debug: bool = False


class DeepSpeedCommsConfig:

    def __init__(self, ds_config):
        self.comms_logger_enabled = COMMS_LOGGER in ds_config
        self.comms_logger = CommsLoggerConfig(**ds_config.get(COMMS_LOGGER, {}))
