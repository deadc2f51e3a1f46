"""`aio` section of the ds_config (reference: runtime/swap_tensor/aio_config.py get_aio_config)."""
AIO_DEFAULT_DICT = {"block_size": 1048576, "queue_depth": 8, "thread_count": 1, "single_submit": False, "overlap_events": True}


def get_aio_config(param_dict):
    out = dict(AIO_DEFAULT_DICT)
    out.update({k: v for k, v in (param_dict.get("aio") or {}).items() if k in AIO_DEFAULT_DICT})
    return out
