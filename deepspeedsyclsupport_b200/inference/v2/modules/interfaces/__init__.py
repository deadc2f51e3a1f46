from .attention_base import DSSelfAttentionBase, DSSelfAttentionRegistry
from .embedding_base import DSEmbeddingBase, DSEmbeddingRegistry
from .linear_base import DSLinearBase, DSLinearRegistry
from .moe_base import DSMoEBase, DSMoERegistry
from .post_norm_base import DSPostNormBase, DSPostNormRegistry
from .pre_norm_base import DSPreNormBase, DSPreNormRegistry
from .unembed_base import DSUnembedBase, DSUnembedRegistry
