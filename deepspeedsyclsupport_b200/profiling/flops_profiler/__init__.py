from .profiler import *  # noqa: F401,F403
