from .bert_sparse_self_attention import BertSparseSelfAttention
from .matmul import MatMul
from .softmax import Softmax
from .sparse_attention_utils import SparseAttentionUtils
from .sparse_self_attention import SparseSelfAttention
from .sparsity_config import (BigBirdSparsityConfig, BSLongformerSparsityConfig, DenseSparsityConfig, FixedSparsityConfig, LocalSlidingWindowSparsityConfig,
                              SparsityConfig, VariableSparsityConfig)
