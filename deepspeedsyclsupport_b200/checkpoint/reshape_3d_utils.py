"""3D (pipe x tensor x data) checkpoint layout descriptor and its contraction onto a smaller layout.
Parity: deepspeed/checkpoint/reshape_3d_utils.py."""
from .reshape_meg_2d import meg_2d_parallel_map, reshape_meg_2d_parallel
from .reshape_utils import get_files, get_files_with_prefix, partition_data
from .constants import LAYER_FILE_PREFIX, MODEL_FILE_PREFIX, ZERO_FILE_PREFIX

PP_DIM = "PP"
TP_DIM = "TP"
DP_DIM = "DP"


class model_3d_desc(object):

    def __init__(self, pp_degree=1, tp_degree=1, dp_degree=1):
        self.pp_degree = pp_degree
        self.tp_degree = tp_degree
        self.dp_degree = dp_degree

    def reshape(self, target_3d_desc, verbose=False):
        """One 2D (pp x tp) map per TARGET dp index, listing the source global ranks folded into each target cell."""
        valid, err = self.can_reshape(target_3d_desc)
        assert valid, ",".join(err)
        tgt_2d_map = reshape_meg_2d_parallel(old_pp_degree=self.pp_degree, old_tp_degree=self.tp_degree, new_pp_degree=target_3d_desc.pp_degree,
                                             new_tp_degree=target_3d_desc.tp_degree, verbose=verbose)
        flat_3d_map = flatten_dp_dimension(meg_2d_map=tgt_2d_map, src_2d_size=self.pp_degree * self.tp_degree, dp_degree=self.dp_degree)
        return unflatten_dp_dimension(meg_2d_map=flat_3d_map, dp_degree=target_3d_desc.dp_degree)

    def get_desc(self):
        return f"{PP_DIM},{TP_DIM},{DP_DIM} = ({self.pp_degree}, {self.tp_degree}, {self.dp_degree})"

    def world_size(self):
        return self.pp_degree * self.tp_degree * self.dp_degree

    def is_valid(self, pp_index, tp_index, dp_index):
        err_msg = []
        for index, degree, dim_name in ((pp_index, self.pp_degree, PP_DIM), (tp_index, self.tp_degree, TP_DIM), (dp_index, self.dp_degree, DP_DIM)):
            if index >= degree:
                err_msg.append(f"{dim_name} indexing error: index {index} >= degree {degree}")
        return len(err_msg) == 0, err_msg

    def can_reshape(self, target_3d_desc):
        err_msg = []
        for tgt, mine, name in ((target_3d_desc.pp_degree, self.pp_degree, PP_DIM), (target_3d_desc.tp_degree, self.tp_degree, TP_DIM),
                                (target_3d_desc.dp_degree, self.dp_degree, DP_DIM)):
            if tgt > mine:
                err_msg.append(f"Expansion reshape not supported - {name}: {mine} ---> {tgt}")
        return len(err_msg) == 0, err_msg


def get_model_3d_descriptor(dir):
    """Infer (pp, tp, dp) from the file names of a checkpoint folder."""
    file_list = get_files(dir)
    zero_file_list = [f for f in file_list if ZERO_FILE_PREFIX in f.split("/")[-1] and f.endswith("_optim_states.pt")]
    num_pp0_files = len(get_files_with_prefix(file_list, f"{LAYER_FILE_PREFIX}01"))
    if num_pp0_files > 0:
        tp_degree = num_pp0_files
        pp_degree = len(get_files_with_prefix(file_list, MODEL_FILE_PREFIX)) // tp_degree
        dp_degree = max(1, len(zero_file_list) // (pp_degree * tp_degree))
    else:
        tp_degree = len(get_files_with_prefix(file_list, MODEL_FILE_PREFIX))
        dp_degree = max(1, len(zero_file_list) // tp_degree) if tp_degree else max(1, len(zero_file_list))
        pp_degree = 0
    return model_3d_desc(pp_degree, tp_degree, dp_degree)


def flatten_dp_dimension(meg_2d_map, src_2d_size, dp_degree):
    new_map = meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree)
    for pp_index in range(meg_2d_map.pp_degree):
        for tp_index in range(meg_2d_map.tp_degree):
            base = meg_2d_map.get_data(pp_index, tp_index)
            for idx in range(dp_degree):
                new_map.add_data(pp_index, tp_index, [r + idx * src_2d_size for r in base])
    return new_map


def unflatten_dp_dimension(meg_2d_map, dp_degree):
    maps = [meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree) for _ in range(dp_degree)]
    for pp_index in range(meg_2d_map.pp_degree):
        for tp_index in range(meg_2d_map.tp_degree):
            for dp_index, ranks in enumerate(partition_data(meg_2d_map.get_data(pp_index, tp_index), dp_degree)):
                maps[dp_index].add_data(pp_index, tp_index, ranks)
    return maps
