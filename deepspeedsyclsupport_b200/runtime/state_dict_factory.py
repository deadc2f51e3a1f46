"""Loading a model-parallel checkpoint at a DIFFERENT tensor-parallel degree: N checkpoint shards are merged when the
job uses fewer TP ranks, or one shard is split when it uses more.

How a tensor is merged / split is decided by a small rule table keyed on substrings of the parameter name
(Megatron-LM naming):  column-parallel weights concatenate along dim 0, row-parallel along dim 1, fused QKV needs the
per-version interleaving handled by merge/split_query_key_value, everything else is replicated.
API parity: deepspeed/runtime/state_dict_factory.py (SDLoaderFactory, SDLoaderBase.load, MegatronSDLoader)."""
import collections
import copy
import json
import os
from abc import ABC, abstractmethod

import torch

from ..utils.logging import logger
from .weight_quantizer import WeightQuantization

AUTO_MODULE_KEY = "auto"

# (substring, action) -- first match wins
_ROW_PARALLEL = ("attention.dense.weight", "mlp.dense_4h_to_h.weight")
_QKV = ("attention.query_key_value", )
_COL_PARALLEL = ("mlp.dense_h_to_4h.weight", "word_embeddings.weight", "mlp.dense_h_to_4h.bias")
_COL_PARALLEL_SPLIT_EXTRA = ("final_linear.weight", )


def _has(key, patterns):
    return any(p in key for p in patterns)


def _cpu(storage, loc):
    return storage


class SDLoaderFactory:

    @staticmethod
    def get_sd_loader_json(json_file, checkpoint_engine):
        """`json_file`: path or dict with "type", "checkpoints", "version" (+ optional "parallelization", "mp_size").  Descriptions
        of bloom / ds_model checkpoints are returned as they are (the inference engine walks them itself)."""
        if isinstance(json_file, dict):
            spec = json_file
        else:
            with open(json_file) as fd:
                spec = json.load(fd)
        kind = spec["type"]
        if kind.lower() in ("bloom", "ds_model"):
            return spec
        return SDLoaderFactory.get_sd_loader(spec["checkpoints"], checkpoint_engine, kind, spec["version"])

    @staticmethod
    def get_sd_loader(ckpt_list, checkpoint_engine, sd_type="Megatron", version=None):
        loaders = {"Megatron": MegatronSDLoader}
        assert sd_type in loaders, "{} checkpoint type is not supported".format(sd_type)
        return loaders[sd_type](ckpt_list, version, checkpoint_engine)


class SDLoaderBase(ABC):
    """Maps (files on disk, requested tensor-parallel degree, rank) to one of three plans:
         same   as many files as ranks     -> read file `rank`
         merge  k files per rank           -> read files [k*rank, k*(rank+1)) and join them   (subclass: merge_state_dict)
         split  k ranks per file           -> read file rank // k, keep slice rank % k        (subclass: split_state_dict)"""

    def __init__(self, ckpt_list, version, checkpoint_engine):
        if checkpoint_engine is None:
            from .checkpoint_engine.torch_checkpoint_engine import TorchCheckpointEngine
            checkpoint_engine = TorchCheckpointEngine()
        self.ckpt_list, self.version, self.checkpoint_engine = ckpt_list, version, checkpoint_engine
        self.module_key = None
        self.check_ckpt_list()

    def _read(self, path):
        return self.checkpoint_engine.load(path, map_location=_cpu)

    def load(self, mp_world_size, mp_rank, module_key=AUTO_MODULE_KEY, is_pipe_parallel=False, quantize=False, quantize_bits=8, quantize_groups=64,
             mlp_extra_grouping=True):
        """-> (path of the file that names this shard, state dict for `mp_rank`, (quantisation scales or None, files merged))."""
        self.module_key = module_key
        files = len(self.ckpt_list)
        first = mp_rank * files // mp_world_size
        if is_pipe_parallel and module_key is not None and mp_world_size != files:
            # pipeline checkpoints hold one mp_rank file per tp rank; the layer files are reshaped elsewhere
            mp_world_size, first = files, 0
        path = self.ckpt_list[first]
        q_args = (quantize, quantize_bits, quantize_groups, mlp_extra_grouping)
        if files > mp_world_size:
            sd, scales, merged = self.merge_state_dict(mp_world_size, mp_rank, *q_args)
            return path, sd, (scales, merged)
        if files < mp_world_size:
            sd, scales = self.split_state_dict(mp_world_size, mp_rank, *q_args)
            return path, sd, (scales, 1)
        assert os.path.exists(path), f"checkpoint shard {path} is missing"
        sd, scales = self._read(path), None
        if quantize:
            quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size)
            module, scales = quantizer.sd_quantize_megatron(self.get_module(sd), quantize_bits, quantize_groups)
            self.set_module(sd, module)
        return path, sd, (scales, 1)

    def get_merge_state_dicts(self, mp_world_size, mp_rank):
        per_rank, rest = divmod(len(self.ckpt_list), mp_world_size)
        assert rest == 0, "Invalid checkpoints and world size for sd merge"
        mine = self.ckpt_list[per_rank * mp_rank:per_rank * (mp_rank + 1)]
        logger.info(f"mp_rank: {mp_rank}, ckpt_list: {mine}")
        return [self._read(path) for path in mine]

    def get_split_state_dict(self, mp_world_size, mp_rank):
        ranks_per_file, rest = divmod(mp_world_size, len(self.ckpt_list))
        assert rest == 0, "Invalid checkpoints and world size for sd split"
        which, offset = divmod(mp_rank, ranks_per_file)
        logger.info(f"mp_rank: {mp_rank}, ckpt_list: {self.ckpt_list[which]}, offset: {offset}")
        return self._read(self.ckpt_list[which]), ranks_per_file, offset

    # ---- where the module's tensors live inside a checkpoint dict: at the top (key None), under a given key, or under
    # ---- whichever of "module" / "model" exists ("auto")
    def _choose_module_key(self, sd):
        found = [k for k in ("module", "model") if k in sd]
        assert len(found) < 2, "checkpoint has both 'model' and 'module' keys, not sure how to proceed"
        assert found, "checkpoint contains neither 'model' or 'module' keys, not sure how to proceed"
        return found[0]

    def _resolved_key(self, sd):
        return self._choose_module_key(sd) if self.module_key == AUTO_MODULE_KEY else self.module_key

    def get_module(self, sd):
        key = self._resolved_key(sd) if self.module_key is not None else None
        return sd if key is None else sd[key]

    def set_module(self, sd, module):
        if self.module_key is None:
            return module
        sd[self._resolved_key(sd)] = module
        return sd

    def check_ckpt_list(self):
        assert len(self.ckpt_list) > 0, "empty checkpoint list"
        head = self._read(self.ckpt_list[0])
        recorded = head.get("mp_world_size") if hasattr(head, "get") else None
        if recorded is not None:  # a DeepSpeed checkpoint knows how many shards it was written with
            assert len(self.ckpt_list) == recorded, f"checkpoint count {len(self.ckpt_list)} is different from saved mp_world_size {recorded}"

    @abstractmethod
    def merge_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def split_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def sanity_check(self, ckpt_file_name):
        ...


class MegatronSDLoader(SDLoaderBase):
    """Fused-QKV layouts by Megatron checkpoint version (np = heads per partition, hn = head size):
         0   : [(3 * np * hn), h]   -> q, k, v are three contiguous thirds of every shard
         1.0 : [(np * hn * 3), h]   and   2.0 : [(np * 3 * hn), h]   -> per-head interleaved: shards simply concatenate"""

    def merge_query_key_value(self, param_list, ckpt_ver):
        if ckpt_ver == 0:
            assert param_list[0].shape[0] % 3 == 0
            third = param_list[0].shape[0] // 3
            thirds = [torch.split(p, third, dim=0) for p in param_list]
            return torch.cat([torch.cat([t[i] for t in thirds], dim=0) for i in range(3)], dim=0)
        if ckpt_ver in (1.0, 2.0):
            return torch.cat(param_list, dim=0)
        assert False, f"checkpoint version: {ckpt_ver} is not supported"

    def split_query_key_value(self, param, num_to_split, offset, ckpt_ver):
        if ckpt_ver == 0:
            assert param.shape[0] % 3 == 0
            q, k, v = torch.split(param, param.shape[0] // 3, dim=0)
            assert q.shape[0] % num_to_split == 0
            size = q.shape[0] // num_to_split
            return torch.cat([torch.split(t, size, dim=0)[offset] for t in (q, k, v)], dim=0)
        if ckpt_ver in (1.0, 2.0):
            assert param.shape[0] % num_to_split == 0
            return torch.split(param, param.shape[0] // num_to_split, dim=0)[offset]
        assert False, f"checkpoint version: {ckpt_ver} is not supported"

    def merge_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64, mlp_extra_grouping=True):
        self.sanity_check(self.ckpt_list[0])
        sd_list = self.get_merge_state_dicts(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd_list[0])
        shards = [self.get_module(sd) for sd in sd_list]
        ckpt_ver = self.get_checkpoint_version(ds_sd)
        logger.info(f"checkpoint version: {ckpt_ver}")
        quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size) if quantize else None
        merged = collections.OrderedDict()
        for key in shards[0].keys():
            values = [sd[key] for sd in shards]
            if _has(key, _ROW_PARALLEL):
                if quantize:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key, merge_dim=1)
                merged[key] = torch.cat(values, dim=1)
            elif _has(key, _QKV):
                if quantize and "attention.query_key_value.weight" in key:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key)
                    merged[key] = torch.cat(values, dim=0)
                elif quantize:
                    merged[key] = torch.cat(values, dim=0)
                else:
                    merged[key] = self.merge_query_key_value(values, ckpt_ver)
            elif _has(key, _COL_PARALLEL):
                if quantize and "mlp.dense_h_to_4h.weight" in key:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key)
                merged[key] = torch.cat(values, dim=0)
            else:
                merged[key] = values[0]
        ds_sd = self.set_module(ds_sd, merged)
        return ds_sd, (quantizer.merge_scales() if quantize else None), len(shards)

    def split_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64, mlp_extra_grouping=True):
        sd, num_to_split, ckpt_offset = self.get_split_state_dict(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd)
        full = self.get_module(sd)
        ckpt_ver = self.get_checkpoint_version(ds_sd)
        logger.info(f"checkpoint version: {ckpt_ver}")
        quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size) if quantize else None
        mine = collections.OrderedDict()
        for key, value in full.items():
            if _has(key, _ROW_PARALLEL):
                assert value.shape[1] % num_to_split == 0
                if quantize:
                    value = quantizer.Quantize([value], quantize_bits, groups, key)[0]
                mine[key] = torch.split(value, value.shape[1] // num_to_split, dim=1)[ckpt_offset]
            elif _has(key, _QKV):
                if quantize and "attention.query_key_value.weight" in key:
                    value = quantizer.Quantize([value], quantize_bits, groups, key)[0]
                mine[key] = self.split_query_key_value(value, num_to_split, ckpt_offset, ckpt_ver)
            elif _has(key, _COL_PARALLEL + _COL_PARALLEL_SPLIT_EXTRA):
                assert value.shape[0] % num_to_split == 0
                if quantize and "mlp.dense_h_to_4h.weight" in key:
                    value = quantizer.Quantize([value], quantize_bits, groups, key)[0]
                mine[key] = torch.split(value, value.shape[0] // num_to_split, dim=0)[ckpt_offset]
            else:
                mine[key] = value
        ds_sd = self.set_module(ds_sd, mine)
        return ds_sd, (quantizer.merge_scales_split(num_to_split) if quantize else None)

    def sanity_check(self, ckpt_file_name):
        required = ["attention.dense.weight", "mlp.dense_4h_to_h.weight", "attention.query_key_value", "mlp.dense_h_to_4h.weight", "mlp.dense_h_to_4h.bias"]
        module = self.get_module(self._read(ckpt_file_name))
        for part in required:
            assert any(part in k for k in module.keys()), f"key: {part} is not found in the checkpoint {ckpt_file_name}"

    def get_checkpoint_version(self, state_dict):
        return self.version if self.version is not None else state_dict.get("checkpoint_version", 0)
