from . import layer  # noqa: F401
