"""(parity: deepspeed/runtime/zero/mics_utils.py:22-203)  Group construction lives in mics.py; this module adds the container the
reference returns and the small helpers around it."""
from dataclasses import dataclass
from typing import List

from torch import Tensor

from ... import comm as dist
from .mics import create_mics_comm_groups as _create_groups


def scale_tensors(tensors: List[Tensor], scale: int):
    for t in tensors:
        t.div_(scale)


@dataclass
class MiCS_CommGroups:
    param_shard_group = None
    param_shard_size = -1
    param_shard_rank = -1
    param_repli_group = None
    param_repli_size = -1
    param_repli_rank = -1
    param_intra_node_group = None
    param_inter_node_shard_group = None


def create_mics_comm_groups(shard_size, dp_group=None, hierarchical_allgather=False, mpu=None) -> MiCS_CommGroups:
    """Shard / replicate groups of this rank as a MiCS_CommGroups.  Hierarchical all-gather groups are not built: inside one
    NVSwitch domain every peer is one hop away, so the flat shard group is already optimal."""
    shard_group, repli_group = _create_groups(shard_size, dp_group, hierarchical_allgather, mpu)
    g = MiCS_CommGroups()
    g.param_shard_group, g.param_repli_group = shard_group, repli_group
    g.param_shard_size, g.param_shard_rank = dist.get_world_size(shard_group), dist.get_rank(shard_group)
    g.param_repli_size, g.param_repli_rank = dist.get_world_size(repli_group), dist.get_rank(repli_group)
    return g


def _generate_mics_config(world_size, ndev_per_node, shard_size, pp_size=1):
    """Rank lists of the shard and replicate groups (pure arithmetic; used by tests and by launch-time validation)."""
    assert world_size % pp_size == 0
    assert (world_size // pp_size) % shard_size == 0, f"dp group size is not dividable by dp_shard_size, (world_size {world_size}, " \
        f"pp_size {pp_size}, dp_shard_size {shard_size})"
    config = {"shard_groups": [], "replicate_groups": [], "span_nodes": max(1, shard_size // ndev_per_node)}
    for base in range(0, world_size, shard_size):
        config["shard_groups"].append(list(range(base, base + shard_size)))
    for i in range(shard_size):
        config["replicate_groups"].append(list(range(i, world_size, shard_size)))
    return config


def _sizes_all_same(groups):
    return all(len(g) == len(groups[0]) for g in groups)
