from .sparsity_config import (SparsityConfig, DenseSparsityConfig, FixedSparsityConfig, VariableSparsityConfig,  # noqa
                              BigBirdSparsityConfig, BSLongformerSparsityConfig, LocalSlidingWindowSparsityConfig)
from .sparse_self_attention import SparseSelfAttention  # noqa: F401
from .bert_sparse_self_attention import BertSparseSelfAttention  # noqa: F401
from .sparse_attention_utils import SparseAttentionUtils  # noqa: F401
