from .top_k_gating import RaggedTopKGating  # noqa: F401
