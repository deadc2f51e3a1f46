"""diffusers AutoencoderKL -> DSVAE, the CUDA-graph replay wrapper around encode / decode (parity:
deepspeed/module_inject/containers/vae.py).  The VAE is wrapped whole: there are no per-layer tensors to extract."""
from ..policy import DSPolicy


def _autoencoder_class():
    """diffusers moved AutoencoderKL between modules across releases; None when diffusers is not installed."""
    try:
        import diffusers
    except ImportError:
        return None
    for path in ("vae", "autoencoder_kl", "autoencoders.autoencoder_kl"):
        mod = diffusers.models
        try:
            for part in path.split("."):
                mod = getattr(mod, part)
            return mod.AutoencoderKL
        except AttributeError:
            continue
    return getattr(diffusers, "AutoencoderKL", None)


class VAEPolicy(DSPolicy):

    def __init__(self):
        super().__init__()
        self._orig_layer_class = _autoencoder_class()

    def match(self, module):
        return self._orig_layer_class is not None and isinstance(module, self._orig_layer_class)

    def match_replaced(self, module):
        from ...model_implementations.diffusers.vae import DSVAE
        return isinstance(module, DSVAE)

    def apply(self, module, enable_cuda_graph=True):
        from ...model_implementations.diffusers.vae import DSVAE
        return DSVAE(module, enable_cuda_graph=enable_cuda_graph)

    def attention(self, client_module):
        return None
