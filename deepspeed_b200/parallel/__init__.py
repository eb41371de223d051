"""Parallelism strategies: AutoTP (tp), AutoEP (auto_ep), sequence parallel lives in ``deepspeed_b200.sequence``."""
from .tp import (AutoTP, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, VocabParallelEmbedding, apply_autotp,
                 shard_tensor)
