"""(parity: inference/v2/model_implementations/mixtral/model.py)  Mistral attention + top-2 mixture of gated-SiLU experts;
the experts run as two grouped GEMMs over tokens sorted by expert."""
from ..inference_transformer_base import DSMoETransformerModelBase


class MixtralInferenceModel(DSMoETransformerModelBase):
    pass
