"""ZeROCheckpoint: the optimizer shard files of a checkpoint folder viewed as a (pp, tp, dp) grid, with lookup of the
files that feed a given target rank after a degree contraction.  Parity: deepspeed/checkpoint/zero_checkpoint.py."""
import torch

from .constants import (BASE_OPTIMIZER_STATE, GROUP_PADDINGS, OPTIMIZER_STATE_DICT, PARTITION_COUNT)
from .reshape_3d_utils import get_model_3d_descriptor, model_3d_desc
from .reshape_utils import basic_folder_validation, get_zero_files, merge_state

GROUP_STATE_KEY = "state"


class ZeROCheckpoint(object):

    def __init__(self, dir):
        basic_folder_validation(dir)
        self.dir = dir
        self.file_list = get_zero_files(dir)
        self.num_files = len(self.file_list)
        assert self.num_files > 0, f"No ZeRO files found in {dir}"
        self.src_3d = get_model_3d_descriptor(dir)
        self.target_3d = model_3d_desc(pp_degree=self.src_3d.pp_degree, tp_degree=self.src_3d.tp_degree, dp_degree=self.src_3d.dp_degree)
        self._3d_file_map = self.src_3d.reshape(self.target_3d)

    def get_src_world_size(self):
        return self.src_3d.world_size()

    def get_src_tp_degree(self):
        return self.src_3d.tp_degree

    def get_src_pp_degree(self):
        return self.src_3d.pp_degree

    def get_src_dp_degree(self):
        return self.src_3d.dp_degree

    def get_file_indices_for_rank(self, pp_index, tp_index, dp_index):
        assert dp_index < len(self._3d_file_map), f"DP index {dp_index} >= DP degree {len(self._3d_file_map)}"
        return self._3d_file_map[dp_index].get_data(pp_index, tp_index)

    def get_files_for_rank(self, pp_index, tp_index, dp_index):
        return [self.file_list[idx] for idx in self.get_file_indices_for_rank(pp_index, tp_index, dp_index)]

    def get_state_for_rank(self, pp_index, tp_index, dp_index, keys_to_ignore=[], strip_tensor_paddings=True):
        merged = None
        for state_file in self.get_files_for_rank(pp_index, tp_index, dp_index):
            sd = torch.load(state_file, map_location=torch.device("cpu"), weights_only=False)
            for key in keys_to_ignore:
                sd.pop(key, None)
            if strip_tensor_paddings:
                self._strip_tensor_paddings(sd)
            merged = sd if merged is None else merge_state(merged, sd)
            self._update_partition_count(merged)
            if strip_tensor_paddings:
                self._clear_group_paddings(merged)
        return merged

    def print_3d_index_map(self, tag=None):
        if tag:
            print(f"3D index map: {tag}")
        for dp_index, _2d_map in enumerate(self._3d_file_map):
            _2d_map.print_data(f"dp = {dp_index}")

    def print_3d_file_map(self, tag=None):
        if tag:
            print(f"3D file map: {tag}")
        for dp_index, _2d_map in enumerate(self._3d_file_map):
            for pp_index in range(_2d_map.pp_degree):
                for tp_index in range(_2d_map.tp_degree):
                    files = [self.file_list[i] for i in _2d_map.get_data(pp_index, tp_index)]
                    print(f"{pp_index}, {tp_index}, {dp_index} => {files}")

    def reshape(self, target_3d_desc: model_3d_desc):
        self.target_3d = target_3d_desc
        self._3d_file_map = self.src_3d.reshape(self.target_3d)

    def _strip_tensor_paddings(self, sd):
        paddings = self._get_optimizer_state(sd, GROUP_PADDINGS)
        state = self._get_optimizer_state(sd, BASE_OPTIMIZER_STATE)
        if not paddings or not isinstance(state, dict):
            return
        for key, pad in enumerate(paddings):
            if pad <= 0 or key not in state:
                continue
            for name, t in state[key].items():
                if torch.is_tensor(t) and t.dim() > 0:
                    state[key][name] = torch.narrow(t, 0, 0, t.numel() - pad).clone()

    def _clear_group_paddings(self, sd):
        paddings = self._get_optimizer_state(sd, GROUP_PADDINGS)
        if paddings:
            sd[OPTIMIZER_STATE_DICT][GROUP_PADDINGS] = [0] * len(paddings)

    def _get_optimizer_state(self, sd, state_key):
        osd = sd.get(OPTIMIZER_STATE_DICT, None)
        return None if osd is None else osd.get(state_key, None)

    def _update_partition_count(self, sd):
        counts = self._get_optimizer_state(sd, PARTITION_COUNT)
        if counts is not None:
            n = self.target_3d.dp_degree
            sd[OPTIMIZER_STATE_DICT][PARTITION_COUNT] = [n] * len(counts) if isinstance(counts, (list, tuple)) else n
