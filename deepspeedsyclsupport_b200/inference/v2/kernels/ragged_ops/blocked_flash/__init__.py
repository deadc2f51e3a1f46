from .blocked_flash import BlockedFlashAttn  # noqa: F401
