"""Pydantic v1 API under either pydantic major version (parity: deepspeed/pydantic_v1.py)."""
try:
    from pydantic.v1 import *  # noqa: F401,F403
except ImportError:
    from pydantic import *  # noqa: F401,F403
