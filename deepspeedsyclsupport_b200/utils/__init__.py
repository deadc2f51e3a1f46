from . import groups  # noqa: F401
from .logging import log_dist, logger  # noqa: F401


def __getattr__(name):
    import importlib
    if name in ("safe_get_full_fp32_param", "safe_set_full_fp32_param", "safe_get_full_grad", "safe_get_full_optimizer_state",
                "safe_set_full_optimizer_state", "safe_get_local_fp32_param", "safe_get_local_grad",
                "safe_get_local_optimizer_state", "safe_set_local_fp32_param", "safe_set_local_optimizer_state"):
        return getattr(importlib.import_module(__name__ + ".tensor_fragment"), name)
    if name == "RepeatingLoader":
        return importlib.import_module(__name__.rsplit(".", 1)[0] + ".runtime.dataloader").RepeatingLoader
    if name in ("OnDevice",):
        return importlib.import_module(__name__ + ".init_on_device").OnDevice
    if name == "instrument_w_nvtx":
        return importlib.import_module(__name__ + ".nvtx").instrument_w_nvtx
    raise AttributeError(name)
