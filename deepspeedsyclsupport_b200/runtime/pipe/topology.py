"""Cartesian process topologies and the pipe x data x model process grid.

Design: a topology is an n-dimensional integer array `ranks` with one named axis per parallelism dimension
(`ranks = arange(world).reshape(dims)`, row-major: the LAST axis varies fastest, so it holds neighbouring GPUs).
Every query of the public API is an index / slice / moveaxis on that array:
    rank of a coordinate      -> ranks[coord]
    coordinate of a rank      -> unravel_index(rank, dims)
    communicator lists of axis-> moveaxis(ranks, axis, -1).reshape(-1, dims[axis])
    filter_match(pipe=0, ...) -> ranks[0, :, ...].ravel()
API parity: deepspeed/runtime/pipe/topology.py:12-456 (ProcessTopology, PipeDataParallelTopology,
PipeModelDataParallelTopology, PipelineParallelGrid and its mpu interface)."""
from collections import namedtuple

import numpy as np

from ... import comm as dist


class ProcessTopology(object):

    def __init__(self, axes, dims):
        assert len(axes) == len(dims)
        self.axes = list(axes)
        self.dims = [int(d) for d in dims]
        self.ProcessCoord = namedtuple("ProcessCoord", self.axes)
        self.ranks = np.arange(int(np.prod(self.dims)), dtype=np.int64).reshape(self.dims)

    # -- coordinate <-> rank
    @property
    def mapping(self):
        return {self.get_coord(int(r)): int(r) for r in self.ranks.ravel()}

    def get_rank(self, **coord_kwargs):
        if set(coord_kwargs) != set(self.axes):
            raise ValueError("get_rank() does not support slices. Use filter_match())")
        idx = tuple(coord_kwargs[a] for a in self.axes)
        for i, d in zip(idx, self.dims):
            assert 0 <= i < d, f"key {coord_kwargs} invalid"
        return int(self.ranks[idx])

    def get_coord(self, rank):
        if not 0 <= rank < self.ranks.size:
            raise ValueError(f"rank {rank} not found in topology.")
        return self.ProcessCoord(*(int(i) for i in np.unravel_index(rank, self.dims)))

    # -- axis queries
    def get_axis_names(self):
        return self.axes

    def get_dim(self, axis):
        return self.dims[self.axes.index(axis)] if axis in self.axes else 0

    def world_size(self):
        return int(self.ranks.size)

    def get_axis_comm_lists(self, axis):
        """Groups of ranks that differ only in their `axis` coordinate (= the communicators along that axis)."""
        if axis not in self.axes:
            return []
        k = self.axes.index(axis)
        return np.moveaxis(self.ranks, k, -1).reshape(-1, self.dims[k]).tolist()

    def get_axis_list(self, axis, idx):
        """All ranks whose `axis` coordinate equals idx."""
        return np.take(self.ranks, idx, axis=self.axes.index(axis)).ravel().tolist()

    def filter_match(self, **filter_kwargs):
        index = tuple(filter_kwargs.get(a, slice(None)) for a in self.axes)
        return np.atleast_1d(self.ranks[index]).ravel().tolist()

    def get_rank_repr(self, rank, omit_axes=["data", "pipe"], inner_sep="_", outer_sep="-"):
        """'model_00'-style tag used in checkpoint file names: every axis not in omit_axes with this rank's coordinate."""
        coord = self.get_coord(rank)
        return outer_sep.join(f"{a}{inner_sep}{getattr(coord, a):02d}" for a in self.axes if a not in frozenset(omit_axes))

    def __str__(self):
        return str(self.mapping)


class PipeDataParallelTopology(ProcessTopology):
    """pipe x data: data-parallel peers are adjacent ranks, so gradient reductions stay on the closest links."""

    def __init__(self, num_pp, num_dp):
        super().__init__(axes=["pipe", "data"], dims=[num_pp, num_dp])


class PipeModelDataParallelTopology(ProcessTopology):
    """pipe x data x model (3D): tensor-parallel peers are adjacent, then data, then pipeline stages."""

    def __init__(self, num_pp, num_mp, num_dp):
        super().__init__(axes=["pipe", "data", "model"], dims=[num_pp, num_dp, num_mp])


def _prime_factors(N):
    if N <= 0:
        raise ValueError("Values must be strictly positive.")
    out, p = [], 2
    while N > 1:
        while N % p == 0:
            out.append(p)
            N //= p
        p += 1
    return out


def _default_topology(world_size):
    """No topology given: alternate the prime factors of the world size between pipe and data."""
    num_pp = num_dp = 1
    for i, prime in enumerate(_prime_factors(world_size)):
        if i % 2 == 0:
            num_pp *= prime
        else:
            num_dp *= prime
    return PipeDataParallelTopology(num_pp=num_pp, num_dp=num_dp)


class PipelineParallelGrid(object):
    """Creates the process groups of a topology and answers the `mpu` queries of the engine.

    Group creation is collective: every rank walks the same lists in the same order and keeps the groups it is in."""

    def __init__(self, topology=None, process_group=None):
        self.global_rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        self._topo = topology if topology is not None else _default_topology(self.world_size)
        topo = self._topo
        self.data_parallel_size = max(topo.get_dim("data"), 1)
        self.pipe_parallel_size = max(topo.get_dim("pipe"), 1)
        self.model_parallel_size = max(topo.get_dim("model"), 1)
        self.slice_parallel_size = self.model_parallel_size
        assert self._is_grid_valid(), "Invalid Grid"
        me = topo.get_coord(self.global_rank)
        self.stage_id = me.pipe
        self.data_parallel_id = me.data
        self.is_first_stage = self.stage_id == 0
        self.is_last_stage = self.stage_id == self.pipe_parallel_size - 1

        # "model" group in the engine's sense: everything that is NOT data parallel (pipe x tensor slices of one replica)
        replica_lists = [sorted(topo.get_axis_list(axis="data", idx=dp)) for dp in range(self.data_parallel_size)]
        mine, self.ds_model_proc_group = self._make_groups(replica_lists)
        self.ds_model_world_size = len(mine)
        self.ds_model_rank = mine.index(self.global_rank)

        self.dp_groups = topo.get_axis_comm_lists("data")
        self.dp_group, self.dp_proc_group = self._make_groups(self.dp_groups)

        self.pipe_groups = topo.get_axis_comm_lists("pipe")
        self.pp_group, self.pp_proc_group = self._make_groups(self.pipe_groups)
        self.p2p_groups = self._build_p2p_groups()

        self.model_groups = topo.get_axis_comm_lists("model")
        self.mp_group = []
        slice_lists = self.model_groups if self.model_parallel_size > 1 else [[r] for r in range(self.world_size)]
        self.slice_group, self.slice_proc_group = self._make_groups(slice_lists)

    def _make_groups(self, rank_lists):
        mine, mine_group = None, None
        for ranks in rank_lists:
            group = dist.new_group(ranks=list(ranks))
            if self.global_rank in ranks:
                mine, mine_group = list(ranks), group
        assert mine_group is not None
        return mine, mine_group

    def _build_p2p_groups(self):
        """[rank, next-stage rank] for every rank (wraps around from the last stage to the first)."""
        nxt = {}
        for stages in self.pipe_groups:
            assert len(stages) == self.pipe_parallel_size
            for i, r in enumerate(stages):
                nxt[r] = stages[(i + 1) % len(stages)]
        assert len(nxt) == self.world_size
        return [[r, nxt[r]] for r in range(self.world_size)]

    def _is_grid_valid(self):
        return self._topo.world_size() == dist.get_world_size()

    def get_stage_id(self):
        return self._topo.get_coord(rank=self.global_rank).pipe

    def get_data_parallel_id(self):
        return self._topo.get_coord(rank=self.global_rank).data

    def stage_to_global(self, stage_id, **kwargs):
        """Global rank of the process at pipeline stage `stage_id` with otherwise my coordinates (overridable)."""
        coord = self._topo.get_coord(self.global_rank)._replace(pipe=stage_id, **kwargs)
        return self._topo.get_rank(**coord._asdict())

    def topology(self):
        return self._topo

    # ---- mpu interface --------------------------------------------------------------------------------------------
    def get_global_rank(self):
        return self.global_rank

    def get_pipe_parallel_rank(self):
        return self.stage_id

    def get_pipe_parallel_world_size(self):
        return self.pipe_parallel_size

    def get_pipe_parallel_group(self):
        return self.pp_proc_group

    def get_data_parallel_rank(self):
        return self.data_parallel_id

    def get_data_parallel_world_size(self):
        return self.data_parallel_size

    def get_data_parallel_group(self):
        return self.dp_proc_group

    def get_model_parallel_rank(self):
        return self.ds_model_rank

    def get_model_parallel_world_size(self):
        return self.ds_model_world_size

    def get_model_parallel_group(self):
        return self.ds_model_proc_group

    def get_slice_parallel_rank(self):
        return self._topo.get_coord(rank=self.global_rank).model if "model" in self._topo.get_axis_names() else 0

    def get_slice_parallel_world_size(self):
        return self.slice_parallel_size

    def get_slice_parallel_group(self):
        return self.slice_proc_group
