"""Key names and defaults of the `aio` config section (parity: deepspeed/runtime/swap_tensor/constants.py), generated from the
table the config reader itself uses (aio_config.AIO_DEFAULT_DICT): AIO_<KEY> is the key string, AIO_<KEY>_DEFAULT its default."""
import json

from .aio_config import AIO_DEFAULT_DICT

AIO = "aio"
for _key, _default in AIO_DEFAULT_DICT.items():
    globals()[f"AIO_{_key.upper()}"] = _key
    globals()[f"AIO_{_key.upper()}_DEFAULT"] = _default
del _key, _default

AIO_FORMAT = '"aio": ' + json.dumps(AIO_DEFAULT_DICT, indent=2)
