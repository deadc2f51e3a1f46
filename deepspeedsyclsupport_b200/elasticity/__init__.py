from .elasticity import compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config  # noqa: F401
from .config import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize  # noqa: F401
from .utils import is_torch_elastic_compatible  # noqa: F401
from .elastic_agent import DSElasticAgent  # noqa: F401,E402
