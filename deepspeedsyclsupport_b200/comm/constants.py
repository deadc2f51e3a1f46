"""Backend names and the `comms_logger` config keys (parity: deepspeed/comm/constants.py)."""
# backend identifiers accepted by init_distributed; on this platform NCCL (GPU) and gloo (host) are live, the others are names
# user configs written for other platforms may carry
for _name in ("nccl", "ccl", "mpi", "gloo", "sccl", "hccl"):
    globals()[f"{_name.upper()}_BACKEND"] = _name
del _name

# AzureML launch conventions
DEFAULT_AML_MASTER_PORT = "54965"
DEFAULT_AML_NCCL_SOCKET_IFNAME = "^docker0,lo"

# comms logger: COMMS_LOGGER_<KEY> is the key string, COMMS_LOGGER_<KEY>_DEFAULT its default
COMMS_LOGGER = "comms_logger"
_LOGGER_DEFAULTS = {"enabled": False, "verbose": False, "prof_all": True, "debug": False, "prof_ops": []}
for _key, _default in _LOGGER_DEFAULTS.items():
    globals()[f"COMMS_LOGGER_{_key.upper()}"] = _key
    globals()[f"COMMS_LOGGER_{_key.upper()}_DEFAULT"] = _default
del _key, _default

COMMS_LOGGER_FORMAT = ('The Comms Logger can be specified as:\n"comms_logger": {"enabled": true, "verbose": false, "prof_all": true, '
                       '"debug": false, "prof_ops": ["all_reduce", "custom_all_reduce_name"]}\n')
