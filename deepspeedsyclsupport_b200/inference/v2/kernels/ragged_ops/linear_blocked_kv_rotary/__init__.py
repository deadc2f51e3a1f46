from .blocked_kv_rotary import BlockedRotaryEmbeddings  # noqa: F401
from .blocked_trained_kv_rotary import BlockedTrainedRotaryEmbeddings  # noqa: F401
from .linear_blocked_kv_copy import LinearBlockedKVCopy  # noqa: F401
