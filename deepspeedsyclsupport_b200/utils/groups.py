"""Process-group factory and getters for data / model(tensor) / sequence / expert / expert-data parallelism and
the ZeRO++ secondary (hpZ) and all-to-all (qgZ) groups.

Parity: deepspeed/utils/groups.py (expert groups :113-293, qgZ a2a groups :356-382, dp getters :385-392,
sequence getters :452-497, hpZ :505-532).  Rank arithmetic is exposed as pure functions so it can be unit-tested
without a process group.  On one 8xB200 NVSwitch box every group spans one uniform fabric, so "intra-node"
hierarchies (hpZ, qgZ two-hop) collapse to the same single hop -- the config keys and getters remain.
"""
from .. import comm as dist
from .logging import log_dist

_WORLD_GROUP = None
_ZERO_PARAM_INTRA_PARALLEL_GROUP = None
_EXPERT_PARALLEL_GROUP = {}
_EXPERT_DATA_PARALLEL_GROUP = {}
_ALL_TO_ALL_GROUP = {}
_DATA_PARALLEL_GROUP = None
_MODEL_PARALLEL_GROUP = None
_SEQUENCE_PARALLEL_GROUP = None
_SEQUENCE_DATA_PARALLEL_GROUP = None
mpu = None
expert_tensor_parallel_world_size = 1


def reset():
    """Forget every cached group (used by tests that rebuild the world at a different size)."""
    global _WORLD_GROUP, _ZERO_PARAM_INTRA_PARALLEL_GROUP, _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP
    global _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP, mpu
    _WORLD_GROUP = _ZERO_PARAM_INTRA_PARALLEL_GROUP = _DATA_PARALLEL_GROUP = _MODEL_PARALLEL_GROUP = None
    _SEQUENCE_PARALLEL_GROUP = _SEQUENCE_DATA_PARALLEL_GROUP = None
    _EXPERT_PARALLEL_GROUP.clear()
    _EXPERT_DATA_PARALLEL_GROUP.clear()
    _ALL_TO_ALL_GROUP.clear()
    mpu = None


def initialize(ep_size=1, mpu=None):
    """Deprecated entry point kept for API parity."""
    raise DeprecationWarning("Please do not use the groups.initialize() API as it is deprecated. "
                             "Instead, pass the desired ep_size to deepspeed.moe.layer.MoE(..,ep_size,..)")


def _ensure_divisibility(numerator, denominator):
    assert numerator % denominator == 0, "{} is not divisible by {}".format(numerator, denominator)


# ----------------------------------------------------------------------------- pure rank arithmetic
def model_parallel_ranks(world_size, mp_size):
    """Consecutive ranks form a model(tensor)-parallel group; returns (mp_groups, dp_groups)."""
    _ensure_divisibility(world_size, mp_size)
    mp_groups = [list(range(i * mp_size, (i + 1) * mp_size)) for i in range(world_size // mp_size)]
    dp_groups = [list(range(i, world_size, mp_size)) for i in range(mp_size)]
    return mp_groups, dp_groups


def expert_parallel_ranks_no_mp(world_size, ep_size, use_data_before_expert_parallel=False):
    """E + D layout (no tensor parallelism).  Returns (expert_parallel_groups, expert_data_parallel_groups)."""
    _ensure_divisibility(world_size, ep_size)
    ep_stride = world_size // ep_size if use_data_before_expert_parallel else 1
    edp_groups, ep_groups = [], []
    for i in range(ep_size):
        if use_data_before_expert_parallel:
            edp_groups.append(list(range(i * ep_stride, (i + 1) * ep_stride)))
        else:
            edp_groups.append(list(range(i, world_size, ep_size)))
    for i in range(world_size // ep_size):
        if use_data_before_expert_parallel:
            ep_groups.append(list(range(i, world_size, ep_stride)))
        else:
            ep_groups.append(list(range(i * ep_size, (i + 1) * ep_size)))
    return ep_groups, edp_groups


def _get_expert_parallel_ranks(world_size, model_parallel_size_, expert_parallel_size_,
                               use_data_before_expert_parallel_=False):
    """E + M + D layout.  Example: world 16, mp 2, ep 4 ->
         expert_parallel_groups      = [0,2,4,6], [8,10,12,14], [1,3,5,7], [9,11,13,15]
         expert_data_parallel_groups = [0,8], [2,10], [4,12], [6,14], [1,9], [3,11], [5,13], [7,15]"""
    _ensure_divisibility(world_size, model_parallel_size_)
    dp_world_size = world_size // model_parallel_size_
    _ensure_divisibility(dp_world_size, expert_parallel_size_)
    data_parallel_groups = [list(range(i, world_size, model_parallel_size_)) for i in range(model_parallel_size_)]
    expert_parallel_groups, expert_data_parallel_groups = [], []
    for dp_ranks in data_parallel_groups:
        part_ep_groups = []
        for i in range(0, dp_world_size, expert_parallel_size_):
            if use_data_before_expert_parallel_:
                stride = dp_world_size // expert_parallel_size_
                part_ep_groups.append([dp_ranks[(i // expert_parallel_size_) + j * stride]
                                       for j in range(expert_parallel_size_)])
            else:
                part_ep_groups.append(dp_ranks[i:i + expert_parallel_size_])
        expert_parallel_groups.extend(part_ep_groups)
        for expert_dp_ranks in zip(*part_ep_groups):
            expert_data_parallel_groups.append(list(expert_dp_ranks))
    return expert_parallel_groups, expert_data_parallel_groups


def sequence_parallel_ranks(world_size, sp_size):
    """Consecutive ranks form a sequence-parallel group; dp groups stride by sp_size; seq x data = world."""
    _ensure_divisibility(world_size, sp_size)
    sp_groups = [list(range(i * sp_size, (i + 1) * sp_size)) for i in range(world_size // sp_size)]
    dp_groups = [list(range(i, world_size, sp_size)) for i in range(sp_size)]
    return sp_groups, dp_groups


# ----------------------------------------------------------------------------- group construction
def _new_groups(rank_lists):
    """Every rank must call new_group for every list; returns the group this rank belongs to."""
    mine = None
    me = dist.get_rank()
    for ranks in rank_lists:
        g = dist.new_group(ranks)
        if me in ranks:
            mine = g
    return mine


def _create_model_parallel(model_parallel_size_):
    """Create (dp_group, mp_group) with consecutive-rank tensor-parallel groups."""
    global _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP
    log_dist(f"Creating model parallel group with size {model_parallel_size_}", ranks=[0])
    mp_groups, dp_groups = model_parallel_ranks(dist.get_world_size(), model_parallel_size_)
    _DATA_PARALLEL_GROUP = _new_groups(dp_groups)
    _MODEL_PARALLEL_GROUP = _new_groups(mp_groups)
    return _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP


def _create_sequence_parallel(sp_size):
    """Ulysses groups: (sp_group, sp_dp_group); ZeRO shards over the full seq x data world."""
    global _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP, _DATA_PARALLEL_GROUP
    sp_groups, dp_groups = sequence_parallel_ranks(dist.get_world_size(), sp_size)
    _SEQUENCE_PARALLEL_GROUP = _new_groups(sp_groups)
    _DATA_PARALLEL_GROUP = _new_groups(dp_groups)
    _SEQUENCE_DATA_PARALLEL_GROUP = dist.get_world_group()
    return _SEQUENCE_PARALLEL_GROUP, _DATA_PARALLEL_GROUP


def _create_expert_and_data_parallel(expert_parallel_size_, use_data_before_expert_parallel_=False):
    assert dist.is_initialized()
    log_dist(f"Creating expert and data parallel groups with size {expert_parallel_size_}", ranks=[0])
    world_size = dist.get_world_size()
    group_name = f"ep_size_{expert_parallel_size_}"
    if group_name in _EXPERT_PARALLEL_GROUP:
        return
    ep_groups, edp_groups = expert_parallel_ranks_no_mp(world_size, expert_parallel_size_,
                                                        use_data_before_expert_parallel_)
    _EXPERT_DATA_PARALLEL_GROUP[group_name] = _new_groups(edp_groups)
    _EXPERT_PARALLEL_GROUP[group_name] = _new_groups(ep_groups)


def _create_expert_data_and_model_parallel(expert_parallel_size_, mpu, use_data_before_expert_parallel_=False):
    assert dist.is_initialized()
    model_parallel_size_ = mpu.get_model_parallel_world_size()
    global expert_tensor_parallel_world_size
    expert_tensor_parallel_world_size = model_parallel_size_
    world_size = dist.get_world_size()
    global _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP
    _DATA_PARALLEL_GROUP = mpu.get_data_parallel_group()
    _MODEL_PARALLEL_GROUP = mpu.get_model_parallel_group()
    group_name = f"ep_size_{expert_parallel_size_}"
    if group_name in _EXPERT_PARALLEL_GROUP:
        return
    ep_groups, edp_groups = _get_expert_parallel_ranks(world_size, model_parallel_size_, expert_parallel_size_,
                                                       use_data_before_expert_parallel_)
    _EXPERT_PARALLEL_GROUP[group_name] = _new_groups(ep_groups)
    _EXPERT_DATA_PARALLEL_GROUP[group_name] = _new_groups(edp_groups)


def _get_max_expert_size():
    assert _EXPERT_PARALLEL_GROUP, "Warning! Process group not initialized"
    return max(int(k.split("_")[2]) for k in _EXPERT_PARALLEL_GROUP)


def _get_max_expert_size_name():
    return f"ep_size_{_get_max_expert_size()}"


def _get_max_expert_parallel_group():
    return _get_expert_parallel_group(_get_max_expert_size_name())


def _get_expert_parallel_group(group_name):
    assert group_name in _EXPERT_PARALLEL_GROUP, "expert parallel group is not initialized"
    return _EXPERT_PARALLEL_GROUP[group_name]


def _get_expert_parallel_group_dict():
    return _EXPERT_PARALLEL_GROUP


def _get_expert_data_parallel_group(group_name):
    assert group_name in _EXPERT_DATA_PARALLEL_GROUP, "expert data parallel group is not initialized"
    return _EXPERT_DATA_PARALLEL_GROUP[group_name]


def _get_expert_data_parallel_group_dict():
    return _EXPERT_DATA_PARALLEL_GROUP


def _clone_world_group():
    global _WORLD_GROUP
    if _WORLD_GROUP is None:
        _WORLD_GROUP = dist.get_world_group()
    return _WORLD_GROUP


def _get_local_all_to_all_group():
    """qgZ groups: 'local_i' (ranks of node i) and 'global_i' (i-th rank of every node).  parity :356-382"""
    assert dist.is_initialized(), "dist is not initialized"
    if _ALL_TO_ALL_GROUP:
        return _ALL_TO_ALL_GROUP
    import os
    world = dist.get_world_size()
    device_per_node = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("LOCAL_SIZE", world)))
    device_per_node = max(1, min(device_per_node, world))
    num_local = world // device_per_node
    if num_local == 0 and world > 0:
        assert world >= 1, "num_gpus must >=1, cannot initialize All-To-All"
        _ALL_TO_ALL_GROUP["local_0"] = dist.new_group(ranks=list(range(world)))
    for i in range(num_local):
        _ALL_TO_ALL_GROUP[f"local_{i}"] = dist.new_group(ranks=[j + device_per_node * i for j in range(device_per_node)])
    if world > device_per_node:
        for i in range(device_per_node):
            _ALL_TO_ALL_GROUP[f"global_{i}"] = dist.new_group(ranks=[i + j * device_per_node for j in range(num_local)])
    return _ALL_TO_ALL_GROUP


# ----------------------------------------------------------------------------- getters
def _get_data_parallel_group():
    assert dist.is_initialized(), "dist is not initialized"
    if mpu is not None:
        return mpu.get_data_parallel_group()
    if _DATA_PARALLEL_GROUP is not None:
        return _DATA_PARALLEL_GROUP
    return _clone_world_group()


def _get_broadcast_src_rank():
    return dist.get_global_rank(_get_sequence_data_parallel_group(), 0)


def _get_expert_broadcast_src_rank(group_name):
    return dist.get_global_rank(_get_expert_data_parallel_group(group_name), 0)


def _get_expert_parallel_world_size(group_name):
    return dist.get_world_size(group=_get_expert_parallel_group(group_name))


def _get_expert_data_parallel_world_size(group_name):
    return dist.get_world_size(group=_get_expert_data_parallel_group(group_name))


def _get_expert_parallel_rank(group_name):
    return dist.get_rank(group=_get_expert_parallel_group(group_name))


def _get_expert_parallel_src_rank(group_name):
    global_rank = dist.get_rank()
    local_world_size = _get_expert_parallel_world_size(group_name)
    return (global_rank // local_world_size) * local_world_size


def _get_expert_data_parallel_rank(group_name):
    return dist.get_rank(group=_get_expert_data_parallel_group(group_name))


def _get_data_parallel_world_size():
    if mpu is not None:
        return mpu.get_data_parallel_world_size()
    return dist.get_world_size(group=_get_data_parallel_group()) if dist.is_initialized() else 1


def _get_model_parallel_group():
    if mpu is not None:
        return mpu.get_model_parallel_group()
    return _MODEL_PARALLEL_GROUP


def _get_model_parallel_world_size():
    if mpu is not None:
        return mpu.get_model_parallel_world_size()
    if _MODEL_PARALLEL_GROUP is not None:
        return dist.get_world_size(group=_MODEL_PARALLEL_GROUP)
    return 1


def _get_model_parallel_rank():
    if mpu is not None:
        return mpu.get_model_parallel_rank()
    if _MODEL_PARALLEL_GROUP is not None:
        return dist.get_rank(group=_MODEL_PARALLEL_GROUP)
    return 0


def _get_data_parallel_rank():
    if mpu is not None:
        return mpu.get_data_parallel_rank()
    return dist.get_rank(group=_get_data_parallel_group()) if dist.is_initialized() else 0


def _get_sequence_parallel_world_size():
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_world_size"):
        return mpu.get_sequence_parallel_world_size()
    if _SEQUENCE_PARALLEL_GROUP is not None:
        return dist.get_world_size(group=_SEQUENCE_PARALLEL_GROUP)
    return 1


def _get_sequence_parallel_rank():
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_rank"):
        return mpu.get_sequence_parallel_rank()
    if _SEQUENCE_PARALLEL_GROUP is not None:
        return dist.get_rank(group=_SEQUENCE_PARALLEL_GROUP)
    return 0


def _get_sequence_parallel_group():
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_group"):
        return mpu.get_sequence_parallel_group()
    return _SEQUENCE_PARALLEL_GROUP


def _get_sequence_data_parallel_world_size():
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_world_size"):
        return mpu.get_sequence_data_parallel_world_size()
    if _SEQUENCE_DATA_PARALLEL_GROUP is not None:
        return dist.get_world_size(group=_SEQUENCE_DATA_PARALLEL_GROUP)
    return _get_data_parallel_world_size()


def _get_sequence_data_parallel_rank():
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_rank"):
        return mpu.get_sequence_data_parallel_rank()
    if _SEQUENCE_DATA_PARALLEL_GROUP is not None:
        return dist.get_rank(group=_SEQUENCE_DATA_PARALLEL_GROUP)
    return _get_data_parallel_rank()


def _get_sequence_data_parallel_group():
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_group"):
        return mpu.get_sequence_data_parallel_group()
    if _SEQUENCE_DATA_PARALLEL_GROUP is not None:
        return _SEQUENCE_DATA_PARALLEL_GROUP
    return _get_data_parallel_group()


def _get_expert_model_parallel_world_size():
    return expert_tensor_parallel_world_size


def _create_zero_param_parallel_group(group_size):
    """hpZ secondary partition group: `group_size` consecutive ranks share a full replica of the shards. parity :505-532"""
    assert dist.is_initialized()
    global _ZERO_PARAM_INTRA_PARALLEL_GROUP
    assert _ZERO_PARAM_INTRA_PARALLEL_GROUP is None, "ZeRO parameter intra parallel group is already initialized"
    world_size = dist.get_world_size()
    zero_param_parallel_size_ = min(group_size, world_size)
    _ensure_divisibility(world_size, zero_param_parallel_size_)
    lists = [list(range(i * zero_param_parallel_size_, (i + 1) * zero_param_parallel_size_))
             for i in range(world_size // zero_param_parallel_size_)]
    _ZERO_PARAM_INTRA_PARALLEL_GROUP = _new_groups(lists)


def _get_zero_param_intra_parallel_group():
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP


def _zero_param_parallel_is_initialized():
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP is not None


def _get_zero_param_intra_parallel_rank_in_mygroup():
    return dist.get_rank(group=_get_zero_param_intra_parallel_group())


def _get_zero_param_intra_parallel_group_world_size():
    return dist.get_world_size(group=_get_zero_param_intra_parallel_group())


def _get_zero_param_intra_parallel_group_ranks():
    return dist.get_all_ranks_from_group(group=_get_zero_param_intra_parallel_group())
