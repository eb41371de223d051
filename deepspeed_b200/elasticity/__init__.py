"""Elastic batch-size math + agent (reference ``deepspeed/elasticity/*``)."""
from .elasticity import (compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config,
                         apply_elasticity_to_config, ElasticityError, ElasticityConfigError,
                         ElasticityIncompatibleWorldSize, get_valid_gpus, get_best_candidates)
