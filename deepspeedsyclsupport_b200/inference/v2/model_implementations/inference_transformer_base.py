"""(parity: inference/v2/model_implementations/inference_transformer_base.py:47-603)  The reference base class builds the
embedding / norm / attention / linear / MoE / unembedding modules from abstract properties of the model (num_layers, model_dim,
head_size, ...) and provides the `transform_*_param` hooks the containers call.  Here that role is played by
`RaggedTransformer`, which takes those facts from an `ArchSpec`; `DSTransformerModelBase` is the same class under the
reference's name plus the property spellings the reference's subclasses and tests use."""
from .inference_model_base import DSInferenceModelBase
from .ragged_transformer import RaggedTransformer


class DSTransformerModelBase(RaggedTransformer, DSInferenceModelBase):
    num_layers = property(lambda self: self.spec.num_layers)
    model_dim = property(lambda self: self.spec.hidden_size)
    vocab_size = property(lambda self: self.spec.vocab_size)
    n_heads = property(lambda self: self.spec.num_heads)
    intermediate_dim = property(lambda self: self.spec.intermediate_size)
    max_sequence_length = property(lambda self: self.spec.max_positions)
    norm_type = property(lambda self: self.spec.norm)
    mlp_activation_fn = property(lambda self: self.spec.activation)
    n_heads_q_local = property(lambda self: self.hq)
    n_heads_kv_local = property(lambda self: self.hkv)
    activation_dtype = property(lambda self: self.dtype)


class DSMoETransformerModelBase(DSTransformerModelBase):
    n_top_k = property(lambda self: self.spec.top_k)
    normalize_expert_scores = property(lambda self: True)
