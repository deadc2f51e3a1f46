"""diffusers AutoencoderKL -> DSVAE (parity: deepspeed/module_inject/containers/vae.py)."""
from ..policy import DSPolicy


class VAEPolicy(DSPolicy):

    def __init__(self):
        super().__init__()
        try:
            import diffusers
            self._orig_layer_class = (diffusers.models.vae.AutoencoderKL if hasattr(diffusers.models, "vae") else
                                      diffusers.models.autoencoder_kl.AutoencoderKL)
        except (ImportError, AttributeError):
            self._orig_layer_class = None

    def match(self, module):
        return self._orig_layer_class is not None and isinstance(module, self._orig_layer_class)

    def match_replaced(self, module):
        from ...model_implementations.diffusers.vae import DSVAE
        return isinstance(module, DSVAE)

    def apply(self, module, enable_cuda_graph=True):
        from ...model_implementations.diffusers.vae import DSVAE
        return DSVAE(module, enable_cuda_graph=enable_cuda_graph)

    def attention(self, client_module):
        """The VAE is wrapped whole; it has no per-layer tensors to extract."""
