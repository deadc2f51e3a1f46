"""Re-mapping a (pipeline x tensor) parallel grid of checkpoint shards onto a smaller grid.

`meg_2d_parallel_map` records, per (pp, tp) cell, the list of source ranks / files whose state must be merged into it;
`reshape_meg_2d_parallel(old_pp, old_tp, new_pp, new_tp)` folds the old grid: the tensor dimension first (consecutive tp
ranks merge), then the pipeline dimension.  Only contraction (new degree divides old degree) is supported, as upstream.
Parity: deepspeed/checkpoint/reshape_meg_2d.py."""
from .reshape_utils import partition_data


class meg_2d_parallel_map(object):

    def __init__(self, pp_degree, tp_degree):
        self.pp_degree = pp_degree
        self.tp_degree = tp_degree
        self.map = {}

    @staticmethod
    def make_key(i, j):
        return f"{i},{j}"

    def _validate_indices(self, pp_index, tp_index):
        assert pp_index < self.pp_degree
        assert tp_index < self.tp_degree

    def simple_init(self):
        self.map = {self.make_key(i // self.tp_degree, i % self.tp_degree): [i] for i in range(self.pp_degree * self.tp_degree)}

    def add_data(self, pp_index, tp_index, data):
        self._validate_indices(pp_index, tp_index)
        assert type(data) is list
        self.map.setdefault(self.make_key(pp_index, tp_index), []).extend(data)

    def get_data(self, pp_index=None, tp_index=None):
        self._validate_indices(pp_index or 0, tp_index or 0)
        pps = list(range(self.pp_degree)) if pp_index is None else [pp_index]
        tps = list(range(self.tp_degree)) if tp_index is None else [tp_index]
        out = []
        for i in pps:
            for j in tps:
                out += self.map[self.make_key(i, j)]
        return out

    def print_data(self, tag):
        print(f"{tag}")
        for key, value in self.map.items():
            print(f"{key} = {value}")


def _reshape_tp_dimension(old_2d_map, new_tp_degree):
    new_map = meg_2d_parallel_map(old_2d_map.pp_degree, new_tp_degree)
    for i in range(old_2d_map.pp_degree):
        for j, ranks in enumerate(partition_data(old_2d_map.get_data(pp_index=i, tp_index=None), new_tp_degree)):
            new_map.add_data(i, j, ranks)
    return new_map


def _reshape_pp_dimension(old_2d_map, new_pp_degree):
    new_map = meg_2d_parallel_map(new_pp_degree, old_2d_map.tp_degree)
    for j in range(old_2d_map.tp_degree):
        for i, ranks in enumerate(partition_data(old_2d_map.get_data(pp_index=None, tp_index=j), new_pp_degree)):
            new_map.add_data(i, j, ranks)
    return new_map


def reshape_meg_2d_parallel(old_pp_degree, old_tp_degree, new_pp_degree, new_tp_degree, verbose=False):
    assert new_pp_degree <= old_pp_degree
    assert new_tp_degree <= old_tp_degree
    current = meg_2d_parallel_map(old_pp_degree, old_tp_degree)
    current.simple_init()
    if verbose:
        current.print_data("original_2d_map:")
    if old_tp_degree != new_tp_degree:
        current = _reshape_tp_dimension(current, new_tp_degree)
        if verbose:
            current.print_data("after_tp_reshape:")
    if old_pp_degree != new_pp_degree:
        current = _reshape_pp_dimension(current, new_pp_degree)
        if verbose:
            current.print_data("final_2d_map:")
    return current


def get_mpu_ranks(tp_size=1, pp_size=1, dp_size=1, virtual_pp_size=None):
    """Rank lists of the tensor / pipeline / data groups of a Megatron-style grid (tp fastest, then dp, then pp)."""
    world_size = tp_size * pp_size * dp_size
    num_pp_groups = world_size // pp_size
    all_dp_group_ranks = []
    for i in range(pp_size):
        start, end = i * num_pp_groups, (i + 1) * num_pp_groups
        for j in range(tp_size):
            all_dp_group_ranks.append(list(range(start + j, end, tp_size)))
    all_tp_group_ranks = [list(range(i * tp_size, (i + 1) * tp_size)) for i in range(world_size // tp_size)]
    all_pp_group_ranks = [list(range(i, world_size, num_pp_groups)) for i in range(num_pp_groups)]
    return all_tp_group_ranks, all_pp_group_ranks, all_dp_group_ranks


def reshape(src, tgt):
    """src / tgt: objects with tp_degree, pp_degree, dp_degree.  Prints the old -> new group mapping (debug aid)."""
    print(f"\n\n*** Reshaping: {src} => {tgt}")
    tp_map = _reshape_tp_dimension_for_sizes(src, tgt)
    print(tp_map)
    return tp_map


def _reshape_tp_dimension_for_sizes(src, tgt):
    m = reshape_meg_2d_parallel(src.pp_degree, src.tp_degree, tgt.pp_degree, tgt.tp_degree)
    return m.map
