from .moe_gather import MoEGather  # noqa: F401
