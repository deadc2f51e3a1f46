"""Elastic batch-size pre-computation: pick one global batch size that is valid for as many GPU counts as possible,
so a job can shrink/grow without changing its effective batch.  Parity: deepspeed/elasticity/elasticity.py:27-378
(v0.1 :83-124, v0.2 with model parallelism :126-189, compute_elastic_config :233-378)."""
import json
import os
from functools import reduce
from math import gcd

from packaging import version as pkg_version

from ..utils.logging import logger
from .config import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize
from .constants import (DEEPSPEED_ELASTICITY_CONFIG, ELASTICITY, ENABLED, ENABLED_DEFAULT, LATEST_ELASTICITY_VERSION,
                        MINIMUM_DEEPSPEED_VERSION)

# highly composite numbers: batch sizes with many divisors
HCN_LIST = [1, 2, 4, 6, 12, 24, 36, 48, 60, 120, 180, 240, 360, 720, 840, 1260, 1680, 2520, 5040, 7560, 10080, 15120,
            20160, 25200, 27720, 45360, 50400, 55440, 83160, 110880, 166320, 221760, 277200, 332640, 498960, 554400,
            665280, 720720]


def _lcm(a, b):
    return a * b // gcd(a, b)


def get_candidate_batch_sizes(base_list, max_acceptable_batch_size):
    candidates = set()
    for base in base_list:
        if base >= max_acceptable_batch_size:
            candidates.add(base)
            continue
        value = max_acceptable_batch_size // base
        index = 0
        while index + 1 < len(HCN_LIST) and HCN_LIST[index + 1] <= value:
            index += 1
        candidates.add(HCN_LIST[index] * base)
    candidates = sorted(candidates)
    logger.debug(f"Candidate batch size: {candidates}")
    return candidates


def get_valid_gpus(batch_size, micro_batches, min_valid_gpus, max_valid_gpus):
    valid = set()
    for micro_batch in micro_batches:
        if batch_size % micro_batch == 0:
            max_gpus = batch_size // micro_batch
            if min_valid_gpus <= max_gpus <= max_valid_gpus:
                valid.add(max_gpus)
            for i in range(1, max_gpus // 2 + 1):
                if i > max_valid_gpus:
                    break
                if i < min_valid_gpus:
                    continue
                if max_gpus % i == 0:
                    valid.add(i)
    return sorted(valid)


def get_best_candidates(candidate_batch_sizes, micro_batches, min_gpus, max_gpus, prefer_larger):
    max_valid, best_valid, final_bs = 0, None, int(min(micro_batches))
    for bs in candidate_batch_sizes:
        cur = get_valid_gpus(bs, micro_batches, min_gpus, max_gpus)
        if len(cur) > max_valid or (len(cur) == max_valid and ((prefer_larger and bs > final_bs) or
                                                                  (not prefer_larger and bs < final_bs))):
            max_valid, best_valid, final_bs = len(cur), cur, bs
    return final_bs, best_valid


def _get_compatible_gpus_v01(micro_batches, max_acceptable_batch_size, min_gpus=None, max_gpus=None, prefer_larger=True):
    min_gpus = min_gpus or 1
    max_gpus = max_gpus or max_acceptable_batch_size // min(micro_batches)
    if not all(mb <= max_acceptable_batch_size for mb in micro_batches):
        raise ValueError(f"All micro batches must be less than or equal to max_acceptable_batch_size: {max_acceptable_batch_size}")
    lcm = reduce(_lcm, micro_batches)
    base_list = list(micro_batches) + [lcm]
    candidates = get_candidate_batch_sizes(base_list, max_acceptable_batch_size)
    return get_best_candidates(candidates, micro_batches, min_gpus, max_gpus, prefer_larger)


def _get_compatible_gpus_v02(micro_batches, max_acceptable_batch_size, current_num_gpus, min_gpus=None, max_gpus=None,
                             prefer_larger=True, num_gpus_per_node=1, model_parallel_size=1):
    if num_gpus_per_node % model_parallel_size != 0:
        raise ElasticityError(f"In Elasticity v0.2, number of GPUs per node:{num_gpus_per_node} should be divisible by "
                              f"model parallel size {model_parallel_size}")

    def get_microbatch(final_batch_size):
        candidate = None
        for mb in micro_batches:
            if final_batch_size // current_num_gpus % mb == 0:
                if candidate is None or (prefer_larger and candidate < mb) or (not prefer_larger and candidate > mb):
                    candidate = mb
        return candidate

    dp_size_per_node = num_gpus_per_node // model_parallel_size
    final_batch_size, valid_world_size = _get_compatible_gpus_v01(
        micro_batches, int(max_acceptable_batch_size / dp_size_per_node), int(min_gpus / num_gpus_per_node),
        int(max_gpus / num_gpus_per_node), prefer_larger=prefer_larger)
    final_batch_size = int(final_batch_size) * dp_size_per_node
    valid_dp_world_size = [i * dp_size_per_node for i in valid_world_size]
    if current_num_gpus // model_parallel_size in valid_dp_world_size:
        return final_batch_size, valid_dp_world_size, get_microbatch(final_batch_size)
    current_dp_size = (current_num_gpus / num_gpus_per_node) * dp_size_per_node
    candidate_batch_sizes = []
    for micro_batch in micro_batches:
        min_batch_size = micro_batch * current_dp_size
        factor = int(max_acceptable_batch_size // min_batch_size)
        candidate_batch_sizes.append(factor * min_batch_size)
    used_microbatch = None
    if prefer_larger:
        candidate_batch_size = max(candidate_batch_sizes)
    else:
        candidate_batch_size = min(candidate_batch_sizes)
    used_microbatch = micro_batches[candidate_batch_sizes.index(candidate_batch_size)]
    return int(candidate_batch_size), [int(current_dp_size)], used_microbatch


def elasticity_enabled(ds_config: dict):
    if ELASTICITY not in ds_config:
        return False
    return ds_config[ELASTICITY].get(ENABLED, ENABLED_DEFAULT)


def ensure_immutable_elastic_config(runtime_elastic_config_dict: dict):
    """The elastic config used at launch (exported via env) must equal the one seen at run time."""
    if DEEPSPEED_ELASTICITY_CONFIG in os.environ:
        scheduler_elastic_config_dict = json.loads(os.environ[DEEPSPEED_ELASTICITY_CONFIG])
        sched = ElasticityConfig(scheduler_elastic_config_dict)
        run = ElasticityConfig(runtime_elastic_config_dict)
        err = "Elastic config '{}={}' seen by resource scheduler does not match config passed to runtime {}={}"
        for attr in ("max_acceptable_batch_size", "micro_batches", "version"):
            if getattr(run, attr) != getattr(sched, attr):
                raise ElasticityConfigError(err.format(attr, getattr(sched, attr), attr, getattr(run, attr)))
    else:
        logger.warning("Unable to find DEEPSPEED_ELASTICITY_CONFIG environment variable, cannot guarantee resource "
                       "scheduler will scale this job using compatible GPU counts.")


def compute_elastic_config(ds_config: dict, target_deepspeed_version: str, world_size=0, return_microbatch=False):
    """Returns (final_batch_size, valid_gpus[, micro_batch_size]) for the `elasticity` block of ds_config."""
    if not isinstance(ds_config, dict):
        raise ValueError("Expected ds_config to be a dictionary but received " f"a {type(ds_config)}, containing: {ds_config}")
    if ELASTICITY not in ds_config:
        raise ElasticityError(f"'{ELASTICITY}' is missing from config json, please add it if running an elastic training job.")
    elastic_config_dict = ds_config[ELASTICITY]
    if not elastic_config_dict.get(ENABLED, ENABLED_DEFAULT):
        raise ElasticityError("Elasticity is disabled, please enable it ('enabled':true) if running an elastic training job.")
    elastic_config = ElasticityConfig(elastic_config_dict)
    model_parallel_size = elastic_config.model_parallel_size
    num_gpus_per_node = elastic_config.num_gpus_per_node
    if model_parallel_size > 1 and float(elastic_config.version) != 0.2:
        raise ElasticityConfigError(f"Elasticity V{elastic_config.version} does not support model-parallel training. "
                                    f"Given model-parallel size: {model_parallel_size}")
    if float(elastic_config.version) > LATEST_ELASTICITY_VERSION:
        raise ElasticityConfigError(f"Attempting to run elasticity version {elastic_config.version} but runtime only "
                                    f"supports up to {LATEST_ELASTICITY_VERSION}")
    base_target = target_deepspeed_version.split("+")[0]
    if pkg_version.parse(base_target) < pkg_version.parse(MINIMUM_DEEPSPEED_VERSION):
        raise ElasticityError(f"Unable to run elasticity on target deepspeed version of {target_deepspeed_version}, "
                              f"currently {MINIMUM_DEEPSPEED_VERSION} is the minimum version supported.")
    if float(elastic_config.version) == 0.1:
        final_batch_size, valid_gpus = _get_compatible_gpus_v01(
            micro_batches=elastic_config.micro_batches, max_acceptable_batch_size=elastic_config.max_acceptable_batch_size,
            min_gpus=elastic_config.min_gpus, max_gpus=elastic_config.max_gpus,
            prefer_larger=elastic_config.prefer_larger_batch_size)
        final_batch_size = int(final_batch_size)
        candidate_microbatch_size = None
    elif float(elastic_config.version) == 0.2:
        if world_size != 0:
            current_num_gpus = world_size
        elif "WORLD_SIZE" in os.environ and os.getenv("WORLD_SIZE").isnumeric():
            current_num_gpus = int(os.getenv("WORLD_SIZE"))
        else:
            raise ElasticityConfigError("Elasticity V 0.2 needs WORLD_SIZE to compute valid batch size. "
                                        "Either give it as argument to function compute_elastic_config or set it as "
                                        "an environment variable. ")
        final_batch_size, valid_gpus, candidate_microbatch_size = _get_compatible_gpus_v02(
            micro_batches=elastic_config.micro_batches, max_acceptable_batch_size=elastic_config.max_acceptable_batch_size,
            current_num_gpus=current_num_gpus, min_gpus=elastic_config.min_gpus, max_gpus=elastic_config.max_gpus,
            prefer_larger=elastic_config.prefer_larger_batch_size, num_gpus_per_node=num_gpus_per_node,
            model_parallel_size=model_parallel_size)
        final_batch_size = int(final_batch_size)
    else:
        raise NotImplementedError(f"Unable to find elastic logic for version: {elastic_config.version}")
    logger.info(f"Valid World Size (GPUs / Model Parallel Size): {valid_gpus}")
    if world_size > 0:
        if world_size not in valid_gpus:
            raise ElasticityIncompatibleWorldSize(f"World size ({world_size}) is not valid with the current list of "
                                                  f"valid GPU counts: {valid_gpus}")
        micro_batch_size = None
        for mbsz in sorted(list(set(elastic_config.micro_batches)), reverse=True):
            if final_batch_size // world_size % mbsz == 0:
                micro_batch_size = mbsz
                break
        assert micro_batch_size is not None, (
            f"Unable to find divisible micro batch size world_size={world_size}, final_batch_size={final_batch_size}, "
            f"and micro_batches={elastic_config.micro_batches}.")
        return final_batch_size, valid_gpus, micro_batch_size
    if return_microbatch:
        if float(elastic_config.version) == 0.2:
            return final_batch_size, valid_gpus, candidate_microbatch_size
        return final_batch_size, valid_gpus, None
    return final_batch_size, valid_gpus
