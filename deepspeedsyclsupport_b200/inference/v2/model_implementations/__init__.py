from .ragged_transformer import ArchSpec, RaggedTransformer, SPEC_FROM_HF  # noqa: F401
