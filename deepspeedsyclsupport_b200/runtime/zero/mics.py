"""MiCS: shard model states inside a sub-group of `mics_shard_size` ranks and replicate across groups; gradients are
reduce-scattered inside the shard group and all-reduced across replica groups.  On one NVSwitch box the hierarchy is
a single hop either way, so this is simply ZeRO-3 with shard group != data-parallel group.
Parity: deepspeed/runtime/zero/mics.py:62-452, mics_utils.py:34-203."""
from ... import comm as dist
from .partition_parameters import Init
from .stage3 import DeepSpeedZeroOptimizer_Stage3


def create_mics_comm_groups(shard_size, dp_group=None, hierarchical_allgather=False, mpu=None):
    """Returns (shard_group, replica_group) for this rank.  ranks [k*shard_size, (k+1)*shard_size) shard together;
    ranks with equal index inside their shard group replicate each other."""
    world = dist.get_world_size(dp_group)
    assert world % shard_size == 0, f"world size {world} not divisible by mics_shard_size {shard_size}"
    ranks = dist.get_all_ranks_from_group(dp_group)
    me = dist.get_rank()
    shard_group = replica_group = None
    for k in range(world // shard_size):
        rl = ranks[k * shard_size:(k + 1) * shard_size]
        g = dist.new_group(rl)
        if me in rl:
            shard_group = g
    for i in range(shard_size):
        rl = ranks[i::shard_size]
        g = dist.new_group(rl)
        if me in rl:
            replica_group = g
    return shard_group, replica_group


class MiCS_AllGatherCoalescedHandle:
    """Completion handle of a shard-group all-gather: ``wait()`` blocks on the collective, then flips every
    covered parameter from in-flight to available (parity: runtime/zero/mics.py:34-59)."""

    def __init__(self, allgather_handle, params, partitions, world_size):
        self.allgather_handle, self.params, self.partitions, self.world_size = allgather_handle, params, partitions, world_size
        self.complete = False

    def wait(self):
        if self.complete:
            return
        from .partition_parameters import ZeroParamStatus
        if self.allgather_handle is not None:
            self.allgather_handle.wait()
        for p in self.params:
            assert p.ds_status == ZeroParamStatus.INFLIGHT, f"expected param {getattr(p, 'ds_id', '?')} to be inflight"
            p.ds_status = ZeroParamStatus.AVAILABLE
        self.complete = True


class MiCS_Init(Init):

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)


class MiCS_Optimizer(DeepSpeedZeroOptimizer_Stage3):

    def __init__(self, module, init_optimizer, *args, mics_shard_size=-1, dp_process_group=None, **kwargs):
        shard_group, replica_group = create_mics_comm_groups(mics_shard_size, dp_process_group)
        super().__init__(module, init_optimizer, *args, dp_process_group=dp_process_group, shard_group=shard_group,
                         replica_group=replica_group, **kwargs)
