from .ragged_embedding import DSRaggedEmbedding
