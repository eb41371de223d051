"""Elastic batch-size math + agent (reference ``deepspeed/elasticity/*``)."""
from .elasticity import (compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config,
                         apply_elasticity_to_config, ElasticityError, ElasticityConfigError,
                         ElasticityIncompatibleWorldSize, get_valid_gpus, get_best_candidates)
from .utils import is_torch_elastic_compatible  # noqa: F401,E402
from .constants import ENABLED, ENABLED_DEFAULT, ELASTICITY  # noqa: F401,E402
if is_torch_elastic_compatible():
    from .elastic_agent import DSElasticAgent  # noqa: F401,E402
