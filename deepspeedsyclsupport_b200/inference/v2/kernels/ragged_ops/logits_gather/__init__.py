from .logits_gather import RaggedLogitsGather  # noqa: F401
