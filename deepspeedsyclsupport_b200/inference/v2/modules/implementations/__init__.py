# importing a sub-package registers its implementations with the interface's registry
from . import attention, embedding, linear, moe, post_norm, pre_norm, unembed
