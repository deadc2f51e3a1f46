from .builder import build_all, build_device, build_host, LIB_DIR  # noqa: F401
