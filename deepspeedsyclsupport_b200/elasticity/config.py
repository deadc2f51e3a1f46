"""`elasticity` config block: validated once, field by field, from a small spec table
(schema: deepspeed/elasticity/config.py; keys/defaults in elasticity/constants.py)."""
import json

from . import constants as K


class ElasticityError(Exception):
    """Base exception for all elasticity related errors."""


class ElasticityConfigError(ElasticityError):
    """Elasticity configuration error."""


class ElasticityIncompatibleWorldSize(ElasticityError):
    """Attempting to run a world size that is incompatible with a given elastic config."""


def _positive(name):

    def check(v):
        if v < 1:
            raise ElasticityConfigError(f"{name} cannot be less than 1, given: {v}")

    return check


class ElasticityConfig(object):
    """
    "elasticity": {"enabled": true, "max_train_batch_size": 2000, "micro_batch_sizes": [2,4,6], "min_gpus": 1,
                   "max_gpus": 10000, "min_time": 20, "prefer_larger_batch": true,
                   "ignore_non_elastic_batch_info": false, "version": 0.1}
    """

    # attribute, key, default, validator
    _OPTIONAL = (
        ("min_gpus", K.MIN_GPUS, K.MIN_GPUS_DEFAULT, None),
        ("max_gpus", K.MAX_GPUS, K.MAX_GPUS_DEFAULT, None),
        ("model_parallel_size", K.MODEL_PARALLEL_SIZE, K.MODEL_PARALLEL_SIZE_DEFAULT, _positive("Model-Parallel size")),
        ("num_gpus_per_node", K.NUM_GPUS_PER_NODE, K.NUM_GPUS_PER_NODE_DEFAULT, _positive("Number of GPUs per node")),
        ("min_time", K.MIN_TIME, K.MIN_TIME_DEFAULT, None),
        ("version", K.VERSION, K.VERSION_DEFAULT, None),
        ("prefer_larger_batch_size", K.PREFER_LARGER_BATCH, K.PREFER_LARGER_BATCH_DEFAULT, None),
        ("ignore_non_elastic_batch_info", K.IGNORE_NON_ELASTIC_BATCH_INFO, K.IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT, None),
    )

    def __init__(self, param_dict):
        self.enabled = param_dict.get(K.ENABLED, K.ENABLED_DEFAULT)
        required = {"max_acceptable_batch_size": (K.MAX_ACCEPTABLE_BATCH_SIZE, K.MAX_ACCEPTABLE_BATCH_SIZE_DEFAULT),
                    "micro_batches": (K.MICRO_BATCHES, K.MICRO_BATCHES_DEFAULT)}
        for attr, (key, default) in required.items():
            if self.enabled and key not in param_dict:
                raise ElasticityConfigError(f"Elasticity config missing {key}")
            setattr(self, attr, param_dict.get(key, default))
        mbs = self.micro_batches
        if not isinstance(mbs, list):
            raise ElasticityConfigError(f"Elasticity expected value of {K.MICRO_BATCHES} to be a list of micro batches, instead is: "
                                        f"{type(mbs)}, containing: {mbs}")
        if any(not isinstance(m, int) for m in mbs):
            raise ElasticityConfigError(f"Elasticity expected {K.MICRO_BATCHES} to only contain a list of integers, instead contains: {mbs}")
        if any(m <= 0 for m in mbs):
            raise ElasticityConfigError(f"Elasticity expected {K.MICRO_BATCHES} to only contain positive integers, instead contains: {mbs}")
        for attr, key, default, validator in self._OPTIONAL:
            value = param_dict.get(key, default)
            if validator is not None:
                validator(value)
            setattr(self, attr, value)
        if min(self.min_gpus, self.max_gpus) < 1:
            raise ElasticityConfigError(f"Elasticity min/max gpus must be > 0, given min_gpus: {self.min_gpus}, max_gpus: {self.max_gpus}")
        if self.max_gpus < self.min_gpus:
            raise ElasticityConfigError(f"Elasticity min_gpus cannot be greater than max_gpus, given min_gpus: {self.min_gpus}, max_gpus: {self.max_gpus}")
        if self.min_time < 0:
            raise ElasticityConfigError(f"Elasticity min time needs to be >= 0: given {self.min_time}")

    def repr(self):
        return dict(self.__dict__)

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4)
