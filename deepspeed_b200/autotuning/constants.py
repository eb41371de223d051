"""Autotuning constants and default tuning spaces (reference ``autotuning/constants.py``)."""
AUTOTUNING = "autotuning"
METRIC_PERCENT_DIFF_CONST = 0.05
DS_CONFIG = "ds_config"
AUTOTUNING_METRIC_LATENCY = "latency"
AUTOTUNING_METRIC_THROUGHPUT = "throughput"
AUTOTUNING_METRIC_FLOPS = "flops"
AUTOTUNING_TUNER_GRIDSEARCH = "gridsearch"
AUTOTUNING_TUNER_RANDOM = "random"
AUTOTUNING_TUNER_MODELBASED = "model_based"
MODEL_INFO = "model_info"
GLOBAL_TUNING_SPACE = "global"
TUNING_MICRO_BATCH_SIZE_PREFIX = "z"

DEFAULT_HF_CONFIG = {
    "train_batch_size": "auto",
    "train_micro_batch_size_per_gpu": "auto",
    "gradient_accumulation_steps": "auto",
}
DEFAULT_MIN_MEM_CONFIG = {"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3},
                          "memory_breakdown": False}
DEFAULT_TUNING_SPACE_ZERO_0 = {"zero_optimization": {"stage": 0}}
DEFAULT_TUNING_SPACE_ZERO_1 = {"zero_optimization": {"stage": 1, "reduce_bucket_size": [5e7, 5e8, 1e9],
                                                     "allgather_bucket_size": [5e7, 5e8, 1e9]}}
DEFAULT_TUNING_SPACE_ZERO_2 = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False],
                                                     "reduce_scatter": [False, True],
                                                     "reduce_bucket_size": [5e7, 5e8, 1e9],
                                                     "allgather_bucket_size": [5e7, 5e8, 1e9],
                                                     "contiguous_gradients": [False, True]}}
DEFAULT_TUNING_SPACE_ZERO_3 = {"zero_optimization": {"stage": 3, "overlap_comm": [True, False],
                                                     "reduce_scatter": [False, True],
                                                     "reduce_bucket_size": [5e7, 5e8, 1e9],
                                                     "allgather_partitions": [True, False],
                                                     "allgather_bucket_size": [5e7, 5e8, 1e9],
                                                     "contiguous_gradients": [False, True]}}
DEFAULT_TUNING_SPACES = {0: DEFAULT_TUNING_SPACE_ZERO_0, 1: DEFAULT_TUNING_SPACE_ZERO_1,
                         2: DEFAULT_TUNING_SPACE_ZERO_2, 3: DEFAULT_TUNING_SPACE_ZERO_3}
