"""Policy plumbing shared by the families: spec translation + model class + container classes + name prefixes."""
from .inference_policy_base import ContainerMap, InferenceV2Policy
from .ragged_transformer import SPEC_FROM_HF


class FamilyPolicy(InferenceV2Policy):
    """Subclasses set: model_types, model_class, transformer_container, non_transformer_container, transformer_prefix,
    unmapped (name suffixes that carry no weights) and, for tied embeddings, tied_source (checkpoint name of the table)."""
    model_class = None
    transformer_container = None
    non_transformer_container = None
    transformer_prefix = "model.layers"
    unmapped = ("rotary_emb.inv_freq", "attn.masked_bias", "attn.bias", "attention.bias", "attention.masked_bias")
    tied_source = None

    def spec(self):
        return SPEC_FROM_HF[self._model_config["model_type"]](self._model_config)

    def instantiate_model(self, engine_config, mp_group, device=None):
        quant = getattr(getattr(engine_config, "quantization", None), "quantization_mode", "")
        return self.model_class(self.spec(), None, tp_group=mp_group, dtype=self._dtype, device=device, quantization_mode=quant,
                                engine_config=engine_config, layer_container_cls=self.transformer_container,
                                non_transformer_container_cls=self.non_transformer_container)

    def build_container_map(self) -> ContainerMap:
        cmap = ContainerMap()
        cmap.set_transformer_params(self.transformer_prefix, self.model.layers)
        cmap.set_non_transformer_params(self.model.non_transformer)
        cmap.set_unmapped_params(self.unmapped)
        self._seen_unembed, self._tied = False, None
        return cmap

    def prepare_tensor(self, name, tensor):
        if name == self.tied_source:
            self._tied = tensor
        if name in ("lm_head.weight", "embed_out.weight"):
            self._seen_unembed = True
        return tensor

    def finish_loading(self, cmap):
        nt = self.model.non_transformer
        if not nt._params["lm_head"].is_finalized and self._tied is not None:  # tied embeddings: no separate lm_head in the file
            nt._params["lm_head"].params = self._tied
