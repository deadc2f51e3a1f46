"""NcclBackend name kept for scripts that import it (parity: deepspeed/runtime/comm/nccl.py:16)."""
from .compressed import CompressedBackend


class NcclBackend(CompressedBackend):
    pass
