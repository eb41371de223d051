"""Typed view of the ``elasticity`` ds_config section (reference ``elasticity/config.py``): validates the fields the
batch-size solver of ``elasticity.py`` relies on and keeps them as attributes."""
import json

from .constants import (ENABLED, ENABLED_DEFAULT, IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT,
                        MAX_ACCEPTABLE_BATCH_SIZE, MAX_ACCEPTABLE_BATCH_SIZE_DEFAULT, MAX_GPUS, MAX_GPUS_DEFAULT,
                        MICRO_BATCHES, MICRO_BATCHES_DEFAULT, MIN_GPUS, MIN_GPUS_DEFAULT, MIN_TIME, MIN_TIME_DEFAULT,
                        MODEL_PARALLEL_SIZE, MODEL_PARALLEL_SIZE_DEFAULT, NUM_GPUS_PER_NODE, NUM_GPUS_PER_NODE_DEFAULT,
                        PREFER_LARGER_BATCH, PREFER_LARGER_BATCH_DEFAULT, VERSION, VERSION_DEFAULT)
from .elasticity import ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize  # noqa: F401


class ElasticityConfig:
    """``ElasticityConfig({"enabled": true, "max_train_batch_size": 2000, "micro_batch_sizes": [2, 4, 6], ...})``.
    An enabled section must name the largest acceptable batch size and the candidate micro-batch sizes."""

    def __init__(self, param_dict):
        self.enabled = param_dict.get(ENABLED, ENABLED_DEFAULT)
        for key, attr, default in ((MAX_ACCEPTABLE_BATCH_SIZE, "max_acceptable_batch_size",
                                    MAX_ACCEPTABLE_BATCH_SIZE_DEFAULT), (MICRO_BATCHES, "micro_batches",
                                                                         MICRO_BATCHES_DEFAULT)):
            if self.enabled and key not in param_dict:
                raise ElasticityConfigError(f"Elasticity config missing {key}")
            setattr(self, attr, param_dict.get(key, default))
        mb = self.micro_batches
        if not isinstance(mb, list):
            raise ElasticityConfigError(f"Elasticity expected value of {MICRO_BATCHES} to be a list of micro batches, "
                                        f"instead is: {type(mb)}, containing: {mb}")
        if not all(isinstance(m, int) and not isinstance(m, bool) for m in mb):
            raise ElasticityConfigError(f"Elasticity expected {MICRO_BATCHES} to only contain a list of integers, "
                                        f"instead contains: {mb}")
        if not all(m > 0 for m in mb):
            raise ElasticityConfigError(f"Elasticity expected {MICRO_BATCHES} to only contain positive integers, "
                                        f"instead contains: {mb}")
        self.min_gpus = param_dict.get(MIN_GPUS, MIN_GPUS_DEFAULT)
        self.max_gpus = param_dict.get(MAX_GPUS, MAX_GPUS_DEFAULT)
        if self.min_gpus < 1 or self.max_gpus < 1:
            raise ElasticityConfigError(f"Elasticity min/max gpus must be > 0, given min_gpus: {self.min_gpus}, "
                                        f"max_gpus: {self.max_gpus}")
        if self.max_gpus < self.min_gpus:
            raise ElasticityConfigError(f"Elasticity min_gpus cannot be greater than max_gpus, given min_gpus: "
                                        f"{self.min_gpus}, max_gpus: {self.max_gpus}")
        self.model_parallel_size = param_dict.get(MODEL_PARALLEL_SIZE, MODEL_PARALLEL_SIZE_DEFAULT)
        if self.model_parallel_size < 1:
            raise ElasticityConfigError(f"Model-Parallel size cannot be less than 1, given model-parallel size: "
                                        f"{self.model_parallel_size}")
        self.num_gpus_per_node = param_dict.get(NUM_GPUS_PER_NODE, NUM_GPUS_PER_NODE_DEFAULT)
        if self.num_gpus_per_node < 1:
            raise ElasticityConfigError(f"Number of GPUs per node cannot be less than 1, given number of GPUs per node: "
                                        f"{self.num_gpus_per_node}")
        self.min_time = param_dict.get(MIN_TIME, MIN_TIME_DEFAULT)
        if self.min_time < 0:
            raise ElasticityConfigError(f"Elasticity min time needs to be >= 0: given {self.min_time}")
        self.version = param_dict.get(VERSION, VERSION_DEFAULT)
        self.prefer_larger_batch_size = param_dict.get(PREFER_LARGER_BATCH, PREFER_LARGER_BATCH_DEFAULT)
        self.ignore_non_elastic_batch_info = param_dict.get(IGNORE_NON_ELASTIC_BATCH_INFO,
                                                            IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT)

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4)
