from .moe_scatter import MoEScatter  # noqa: F401
