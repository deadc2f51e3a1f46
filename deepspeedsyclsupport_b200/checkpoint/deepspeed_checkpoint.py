"""Read-only view of a Megatron-DeepSpeed style checkpoint folder, optionally contracted to smaller parallel degrees.

A folder holds three families of files: `layer_XX-model_YY-model_states.pt` (one per layer per tp rank), `mp_rank_XX_model_states.pt`
(one per (pp, tp) cell) and the ZeRO optimizer shards.  `DeepSpeedCheckpoint(dir, tp, pp, dp)` answers, for a TARGET grid with
degrees <= the source's: which files feed cell (pp, tp[, dp]) and what their merged state is.  Everything is derived from three
small tables built once in `__init__`:

    layer_keys                sorted "layer_XX" prefixes; first = embedding, last = final norm, the rest transformer layers
    _cell_layers[pp]          the transformer layer keys owned by target pipeline stage pp
    _layer_tp_files[key][tp]  the source files of layer `key` folded into target tp rank `tp`

Method names follow deepspeed/checkpoint/deepspeed_checkpoint.py (the universal-checkpoint tools call them).
"""
import os
from functools import lru_cache
from typing import Dict

import torch

from .constants import LAYER_FILE_PREFIX, MODEL_FILE_PREFIX
from .reshape_3d_utils import model_3d_desc
from .reshape_meg_2d import reshape_meg_2d_parallel
from .reshape_utils import basic_folder_validation, get_files, get_files_with_prefix, merge_state, partition_data
from .zero_checkpoint import ZeROCheckpoint

EMBEDDING_LAYER_INDEX = 0
FINAL_LAYER_NORM_INDEX = -1
ARGS_KEY = "args"
CHECKPOINT_INFO_KEY = "checkpoint_info"
ITERATION_KEY = "iteration"

# tensor-parallel merge rules for per-layer files: replicated tensors are taken from the first tp rank, row-parallel weights
# are concatenated along dim 1, everything else (column-parallel) along dim 0
SEQUENTIAL_LAYERS = ["input_layernorm.weight", "input_layernorm.bias", "self_attention.dense.bias", "post_attention_layernorm.weight",
                     "post_attention_layernorm.bias", "mlp.dense_4h_to_h.bias", "position_embeddings.weight"]
LAYER_CONCAT_DIM = {"self_attention.dense.weight": 1, "mlp.dense_4h_to_h.weight": 1}


@lru_cache(maxsize=8)
def _read(path):
    return torch.load(path, map_location=torch.device("cpu"), weights_only=False)


def _load(path):
    """Fresh (uncached) read: merged results may be mutated by callers."""
    return torch.load(path, map_location=torch.device("cpu"), weights_only=False)


class DeepSpeedCheckpoint(object):

    def __init__(self, dir, tp_degree=None, pp_degree=None, dp_degree=None):
        basic_folder_validation(dir)
        self.dir = dir
        self.file_list = get_files(dir)
        self.layer_files = get_files_with_prefix(self.file_list, LAYER_FILE_PREFIX)
        self.mp_rank_files = get_files_with_prefix(self.file_list, MODEL_FILE_PREFIX)
        self.zero_checkpoint = zc = ZeROCheckpoint(dir)
        src = (zc.get_src_pp_degree(), zc.get_src_tp_degree(), zc.get_src_dp_degree())
        self.pp_degree, self.tp_degree, self.dp_degree = (s if want is None else want for s, want in zip(src, (pp_degree, tp_degree, dp_degree)))
        self.original_world_size = zc.get_src_world_size()
        self.world_size = self.tp_degree * max(self.pp_degree, 1) * self.dp_degree
        prefix_len = len(LAYER_FILE_PREFIX) + 2
        self.layer_keys = sorted({os.path.basename(f)[:prefix_len] for f in self.layer_files})
        self.layer_count = len(self.layer_keys)
        self.old_2d_map = self.new_2d_map = None
        self.global_state = {}
        if (self.pp_degree, self.tp_degree, self.dp_degree) != src:
            zc.reshape(model_3d_desc(self.pp_degree, self.tp_degree, self.dp_degree))
        if self.pp_degree > 0 and self.layer_count > 0:
            self.old_2d_map = reshape_meg_2d_parallel(src[0], src[1], src[0], src[1])
            self.new_2d_map = reshape_meg_2d_parallel(src[0], src[1], self.pp_degree, self.tp_degree)
            self._sanity_check()
            self._layer_tp_files = {key: partition_data(get_files_with_prefix(self.layer_files, key), self.tp_degree) for key in self.layer_keys}
            body = self.layer_keys[1:-1]
            self._cell_layers = partition_data(body, self.pp_degree)
            # the tables under the reference's names
            self.pp_to_transformer_map = dict(enumerate(self._cell_layers))
            self.tp_to_embedding_map = dict(enumerate(self._layer_tp_files[self.layer_keys[EMBEDDING_LAYER_INDEX]]))
            self.tp_to_final_norm_map = dict(enumerate(self._layer_tp_files[self.layer_keys[FINAL_LAYER_NORM_INDEX]]))
            self.transformer_file_map = {(tp, pp): [self._layer_tp_files[key][tp] for key in keys]
                                         for pp, keys in enumerate(self._cell_layers) for tp in range(self.tp_degree)}
            first = _read(self.mp_rank_files[0])
            self.global_state.update({ITERATION_KEY: first.get(ITERATION_KEY, 0), ARGS_KEY: first.get(ARGS_KEY, None)})

    # ---- degrees
    def is_change_tp_degree(self):
        return self.tp_degree != self.zero_checkpoint.get_src_tp_degree()

    def is_change_pp_degree(self):
        return self.pp_degree != self.zero_checkpoint.get_src_pp_degree()

    def is_change_dp_degree(self):
        return self.dp_degree != self.zero_checkpoint.get_src_dp_degree()

    def _sanity_check(self):
        cells = self.pp_degree * self.tp_degree
        assert len(self.mp_rank_files) % self.tp_degree == 0, "model-state files do not split evenly over the tp degree"
        assert self.zero_checkpoint.num_files % cells == 0, "optimizer shard files do not split evenly over the (pp, tp) grid"
        assert self.layer_count > 2 and (self.layer_count - 2) % self.pp_degree == 0, "transformer layers do not split evenly over pp"

    def _check_cell(self, tp_index, pp_index=0):
        assert 0 <= tp_index < self.tp_degree and 0 <= pp_index < self.pp_degree, f"cell (pp={pp_index}, tp={tp_index}) outside the target grid"

    # ---- global entries of the first model-state file
    def _get_checkpoint_value(self, key):
        if key not in self.global_state:
            self.global_state[key] = _read(self.mp_rank_files[0]).get(key, None)
        return self.global_state[key]

    def get_iteration(self):
        value = self._get_checkpoint_value(ITERATION_KEY)
        return 0 if value is None else value

    def get_args(self):
        return self._get_checkpoint_value(ARGS_KEY)

    def get_checkpoint_info(self, info_key=CHECKPOINT_INFO_KEY):
        return self._get_checkpoint_value(info_key)

    # ---- ZeRO shards
    def get_zero_checkpoint_state(self, pp_index, tp_index, dp_index) -> dict:
        return self.zero_checkpoint.get_state_for_rank(pp_index=pp_index, tp_index=tp_index, dp_index=dp_index,
                                                       keys_to_ignore=[CHECKPOINT_INFO_KEY])

    def get_zero_files(self, pp_index, tp_index, dp_index) -> list:
        return self.zero_checkpoint.get_files_for_rank(pp_index=pp_index, tp_index=tp_index, dp_index=dp_index)

    # ---- per-layer files
    def _merge_state_dicts(self, sd_list):
        head = sd_list[0]
        return {name: head[name] if name in SEQUENTIAL_LAYERS else torch.cat([sd[name] for sd in sd_list], dim=LAYER_CONCAT_DIM.get(name, 0))
                for name in head}

    def _merged_layer_state(self, files):
        return self._merge_state_dicts([_load(f) for f in files])

    def get_embedding_layer_id(self):
        return self.layer_keys[EMBEDDING_LAYER_INDEX]

    def get_final_norm_layer_id(self):
        return self.layer_keys[FINAL_LAYER_NORM_INDEX]

    def get_embedding_files(self, tp_index: int) -> list:
        self._check_cell(tp_index)
        return self.tp_to_embedding_map[tp_index]

    def get_embedding_state(self, tp_index: int) -> Dict:
        return self._merged_layer_state(self.get_embedding_files(tp_index))

    def get_final_norm_files(self, tp_index: int) -> list:
        self._check_cell(tp_index)
        return self.tp_to_final_norm_map[tp_index]

    def get_final_norm_state(self, tp_index: int) -> Dict:
        return _load(self.get_final_norm_files(tp_index)[0])  # replicated across tp: any source rank's copy

    def get_pp_transformer_map(self, pp_index: int) -> list:
        self._check_cell(0, pp_index)
        return self.pp_to_transformer_map[pp_index]

    def get_transformer_state(self, tp_index: int, pp_index: int) -> list:
        self._check_cell(tp_index, pp_index)
        return [self._merged_layer_state(files) for files in self.transformer_file_map[(tp_index, pp_index)]]

    # ---- mp_rank files
    def get_2d_parallel_files(self, tp_index: int, pp_index: int) -> list:
        self._check_cell(tp_index, pp_index)
        return [self.mp_rank_files[i] for i in self.new_2d_map.get_data(pp_index=pp_index, tp_index=tp_index)]

    def get_2d_parallel_state(self, tp_index: int, pp_index: int) -> dict:
        merged = None
        for path in self.get_2d_parallel_files(tp_index, pp_index):
            sd = _load(path)
            merged = sd if merged is None else merge_state(merged, sd)
        return merged

    # ---- diagnostics
    def validate_files(self):
        for path in self.file_list:
            if not os.path.isfile(path):
                print(f"Error: {path} is not existent")

    def _dump_mapping(self, data_map, map_tag=None):
        if map_tag is not None:
            print(f"Dump mapping: {map_tag}")
        for key, value in data_map.items():
            print(f"{key} = {value}")

    def show_2d_mapping(self):
        print("reshaped 2d map ---- begin")
        for pp in range(self.pp_degree):
            for tp in range(self.tp_degree):
                print(f"[{pp}, {tp}] = {self.get_2d_parallel_files(pp_index=pp, tp_index=tp)}")
        print("reshaped 2d map ---- end")

    def show_tp_embedding_map(self):
        self._dump_mapping(self.tp_to_embedding_map, "tp_to_embedding_layers")

    def show_tp_final_norm_map(self):
        self._dump_mapping(self.tp_to_final_norm_map, "tp_to_final_norm_layers")

    def show_pp_transformer_map(self):
        self._dump_mapping(self.pp_to_transformer_map, "pp_to_transformer_layers")

    def show_transformer_file_map(self):
        self._dump_mapping(self.transformer_file_map, "rank_to_transformer_files")
