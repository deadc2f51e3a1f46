from .ragged_unembed import DSRaggedUnembed
