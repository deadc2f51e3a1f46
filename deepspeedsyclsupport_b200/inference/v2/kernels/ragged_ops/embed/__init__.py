from .embed import RaggedEmbeddingKernel  # noqa: F401
