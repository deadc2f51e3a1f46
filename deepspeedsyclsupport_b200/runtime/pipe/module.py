"""PipelineModule: a model given as a flat sequence of layers, cut into contiguous stage slices.

Pieces (public API parity: deepspeed/runtime/pipe/module.py:86-636):
  LayerSpec / TiedLayerSpec   deferred construction -- a stage only ever instantiates the layers it owns
  stage assignment            `uniform` (layer count), `parameters` (trainable parameter count), `type:<regex>` (count of
                              layers whose class name matches); a prefix-sum balanced split in all three cases
  forward                     the stage's layers run in segments of `activation_checkpoint_interval` layers, each
                              segment optionally under activation checkpointing
  tied layers                 one module per tie key per stage; the stages sharing a key form a process group used to
                              broadcast the initial weight and to all-reduce the weight gradient every step
  checkpoints                 one file per layer (`layer_XX-model_YY-model_states.pt`), so a checkpoint can be reloaded
                              under a different stage assignment
"""
import glob
import os
import re
from functools import partial

import torch
import torch.nn as nn

from ... import comm as dist
from ...accelerator import get_accelerator
from ...utils.logging import logger
from .. import utils as ds_utils
from ..activation_checkpointing import checkpointing
from .topology import PipeDataParallelTopology, PipelineParallelGrid


class PipelineError(Exception):
    """Errors related to the use of deepspeed.PipelineModule"""


class LayerSpec(object):
    """Recipe for a layer: LayerSpec(nn.Linear, 16, 16) builds nn.Linear(16, 16) on the owning stage only."""

    def __init__(self, typename, *module_args, **module_kwargs):
        if not issubclass(typename, nn.Module):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.typename = typename
        self.module_args = module_args
        self.module_kwargs = module_kwargs
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        return ds_utils.call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log=False):
        if log:
            logger.info(f"RANK={self.global_rank} building {repr(self)}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    """A LayerSpec whose `tied_weight_attr` parameter is shared by every layer with the same `key` (e.g. embeddings)."""

    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr="weight", **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key = key
        self.forward_fn = forward_fn
        self.tied_weight_attr = tied_weight_attr


def _display_name(layer):
    if isinstance(layer, LayerSpec):
        return layer.typename.__name__
    if isinstance(layer, nn.Module):
        return type(layer).__name__
    return getattr(layer, "__name__", None)


def _trainable_numel(layer):
    if isinstance(layer, LayerSpec):
        layer = layer.build()
    if isinstance(layer, nn.Module):
        return sum(p.numel() for p in layer.parameters() if p.requires_grad)
    return 0


def _stage_boundaries(specs, method, num_stages):
    """List of num_stages + 1 layer indices; stage s owns [parts[s], parts[s+1])."""
    method = method.lower()
    if method == "uniform":
        return ds_utils.partition_uniform(num_items=len(specs), num_parts=num_stages)
    if method == "parameters":
        return ds_utils.partition_balanced(weights=[_trainable_numel(l) for l in specs], num_parts=num_stages)
    if method.startswith("type:"):
        pattern = re.compile(method.split(":", 1)[1], re.IGNORECASE)
        hits = [1 if (_display_name(l) is not None and pattern.search(_display_name(l))) else 0 for l in specs]
        if not any(hits):
            raise RuntimeError(f"Partitioning '{method.split(':', 1)[1]}' found no valid layers to partition.")
        return ds_utils.partition_balanced(weights=hits, num_parts=num_stages)
    raise NotImplementedError(f"Partitioning method {method} not implemented.")


class PipelineModule(nn.Module):
    checkpoint_parallel_write_pipeline = False

    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seed_layers=False, seed_fn=None, base_seed=1234,
                 partition_method="parameters", activation_checkpoint_interval=0, activation_checkpoint_func=checkpointing.checkpoint,
                 checkpointable_layers=None):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        if checkpointable_layers is not None:
            assert isinstance(checkpointable_layers, list), "param `checkpointable_layers` must be type of list."
        self.loss_fn = loss_fn
        self.checkpointable_layers = checkpointable_layers
        self.seed_layers, self.seed_fn, self.base_seed = seed_layers, seed_fn, base_seed
        self.activation_checkpoint_interval = activation_checkpoint_interval
        self.activation_checkpoint_func = activation_checkpoint_func
        self.micro_offset = 0

        self.world_group = dist.new_group(ranks=range(dist.get_world_size()))
        self.global_rank = dist.get_rank(group=self.world_group)
        self.world_size = dist.get_world_size(group=self.world_group)
        self.local_rank = int(os.environ.get("LOCAL_RANK") or 0)
        if topology is None:
            if self.world_size % num_stages != 0:
                raise RuntimeError(f"num_stages ({num_stages}) must divide distributed world size ({self.world_size})")
            topology = PipeDataParallelTopology(num_pp=num_stages, num_dp=self.world_size // num_stages)
        self._topo = topology
        self.num_stages = topology.get_dim("pipe")
        self._grid = PipelineParallelGrid(process_group=self.world_group, topology=topology)
        self.stage_id = topology.get_coord(self.global_rank).pipe

        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self._partition_layers(method=partition_method)

        self.forward_funcs = []
        self.fwd_map = {}
        self.tied_modules = nn.ModuleDict()
        self.tied_weight_attrs = {}
        self._build()
        on_gpu = get_accelerator().device_name() == "cuda"
        self.to(get_accelerator().device_name(self.local_rank) if on_gpu else "cpu")
        self.tied_comms = self._index_tied_modules()
        self._synchronize_tied_weights()

    # ------------------------------------------------------------------ construction
    def _seed_for(self, value):
        (self.seed_fn or ds_utils.set_random_seed)(value)

    def _register_layer(self, layer_idx, module):
        self.fwd_map[str(layer_idx)] = len(self.forward_funcs)
        self.forward_funcs.append(module)
        self.add_module(str(layer_idx), module)

    def _build(self):
        for layer_idx in range(self._local_start, self._local_stop):
            spec = self._layer_specs[layer_idx]
            if self.seed_layers:
                self._seed_for(self.base_seed + layer_idx)
            if isinstance(spec, PipelineModule):
                raise NotImplementedError("RECURSIVE BUILD NOT YET IMPLEMENTED")
            if isinstance(spec, TiedLayerSpec):
                if spec.key not in self.tied_modules:
                    self.tied_modules[spec.key] = spec.build()
                    self.tied_weight_attrs[spec.key] = spec.tied_weight_attr
                owner = self.tied_modules[spec.key]
                self.forward_funcs.append(owner if spec.forward_fn is None else partial(spec.forward_fn, owner))
            elif isinstance(spec, LayerSpec):
                self._register_layer(layer_idx, spec.build())
            elif isinstance(spec, nn.Module):
                self._register_layer(layer_idx, spec)
            else:
                self.forward_funcs.append(spec)  # plain callable (reshape lambdas etc.)
        for p in self.parameters():
            p.ds_pipe_replicated = False

    def _count_layer_params(self):
        return [_trainable_numel(l) for l in self._layer_specs]

    def _find_layer_type(self, layername):
        pattern = re.compile(layername, re.IGNORECASE)
        idxs = [i for i, l in enumerate(self._layer_specs) if _display_name(l) is not None and pattern.search(_display_name(l))]
        if not idxs:
            raise RuntimeError(f"Partitioning '{layername}' found no valid layers to partition.")
        return idxs

    def _partition_layers(self, method="uniform"):
        num_stages = self._topo.get_dim("pipe")
        if self.global_rank == 0:
            logger.info(f"Partitioning pipeline stages with method {method}")
        self.parts = _stage_boundaries(self._layer_specs, method, num_stages)
        if self.global_rank == 0:
            for stage in range(num_stages):
                lo, hi = self.parts[stage], self.parts[stage + 1]
                print(f"stage={stage} layers={hi - lo}")
                for idx in range(lo, hi):
                    print(f"    {idx:2d}: {_display_name(self._layer_specs[idx]) or str(self._layer_specs[idx])}")
            if self.loss_fn:
                print(f"  loss: {getattr(self.loss_fn, '__name__', type(self.loss_fn).__name__)}")
        self._set_bounds(start=self.parts[self.stage_id], stop=self.parts[self.stage_id + 1])

    def _set_bounds(self, start=None, stop=None):
        self._local_start, self._local_stop = start, stop

    # ------------------------------------------------------------------ forward
    def _segment_fn(self, start, end):
        """Callable running forward_funcs[start:end]; reseeds per layer when seed_layers is on (so recomputation under
        activation checkpointing sees the same dropout masks)."""
        micro = self.micro_offset + 1

        def run(*inputs):
            x = inputs[0] if len(inputs) == 1 else inputs
            for k in range(start, end):
                self.curr_layer = k + self._local_start
                if self.seed_layers:
                    self._seed_for(self.base_seed * micro + self.curr_layer)
                x = self.forward_funcs[k](x)
            return x

        return run

    def forward(self, forward_input):
        self.micro_offset += 1
        n = len(self.forward_funcs)
        interval = self.activation_checkpoint_interval
        if interval == 0:
            return self._segment_fn(0, n)(forward_input)
        x = forward_input
        for start in range(0, n, interval):
            end = min(start + interval, n)
            args = x if isinstance(x, tuple) else (x, )
            seg = self._segment_fn(start, end)
            x = self.activation_checkpoint_func(seg, *args) if self._is_checkpointable(self.forward_funcs[start:end]) else seg(*args)
        return x

    def _is_checkpointable(self, funcs):
        if type(self).__name__ in ("GPTModelPipe", "GPT2ModelPipe"):  # Megatron convention: only the transformer blocks
            return all("ParallelTransformerLayerPipe" in type(f).__name__ for f in funcs)
        if self.checkpointable_layers is not None:
            return all(type(f).__name__ in self.checkpointable_layers for f in funcs)
        return any(isinstance(f, nn.Module) and next(f.parameters(), None) is not None for f in funcs)

    # ------------------------------------------------------------------ tied weights
    def _index_tied_modules(self):
        """{key: {ranks, group, weight_attr, module}} for the tie groups this rank belongs to.  Group creation is
        collective, so every rank creates every group (same order) and keeps its own."""
        comms = {}
        if self._topo.get_dim("pipe") == 1:
            return comms
        keys = []
        for s in self._layer_specs:
            if isinstance(s, TiedLayerSpec) and s.key not in keys:
                keys.append(s.key)
        slices = self._grid.get_slice_parallel_world_size()
        for key in keys:
            stages = sorted({self.stage_owner(i) for i, s in enumerate(self._layer_specs) if isinstance(s, TiedLayerSpec) and s.key == key})
            for dp in range(self._grid.data_parallel_size):
                for mp in range(slices):
                    extra = {"model": mp} if slices > 1 else {}
                    ranks = [self._grid.stage_to_global(stage_id=s, data=dp, **extra) for s in stages]
                    group = dist.new_group(ranks=ranks)
                    if self.global_rank not in ranks:
                        continue
                    assert key in self.tied_modules
                    comms[key] = {"ranks": ranks, "group": group, "weight_attr": self.tied_weight_attrs[key], "module": self.tied_modules[key]}
                    if self.global_rank != ranks[0]:
                        # only the first owner counts the tied parameters (norms, checkpoints)
                        for p in self.tied_modules[key].parameters():
                            p.ds_pipe_replicated = True
        return comms

    def _tied_weight(self, key):
        return getattr(self.tied_modules[key], self.tied_comms[key]["weight_attr"])

    def _synchronize_tied_weights(self):
        for key, comm in self.tied_comms.items():
            dist.broadcast(self._tied_weight(key).data, src=min(comm["ranks"]), group=comm["group"])

    def allreduce_tied_weight_gradients(self):
        for key, comm in self.tied_comms.items():
            w = self._tied_weight(key)
            grad = w.grad if w.grad is not None else getattr(w, "ds_tied_grad", None)
            if grad is not None:
                dist.all_reduce(grad, group=comm["group"])

    def get_tied_weights_and_groups(self):
        return [(self._tied_weight(key), comm["group"]) for key, comm in self.tied_comms.items()]

    # ------------------------------------------------------------------ introspection
    def partitions(self):
        return self.parts

    def stage_owner(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers
        for stage in range(self._topo.get_dim("pipe")):
            if self.parts[stage] <= layer_idx < self.parts[stage + 1]:
                return stage
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.parts}")

    def set_checkpoint_interval(self, interval):
        assert interval >= 0
        self.checkpoint_interval = interval

    def topology(self):
        return self._topo

    def mpu(self):
        return self._grid

    def num_pipeline_stages(self):
        return self._topo.get_dim("pipe")

    # ------------------------------------------------------------------ per-layer checkpoint files
    def ckpt_prefix(self, checkpoints_path, tag):
        coord = self._topo.get_coord(rank=self.global_rank)
        tags = [f"{axis}_{getattr(coord, axis):02d}" for axis in self._topo.get_axis_names() if axis != "data"]
        return os.path.join(checkpoints_path, str(tag), "-".join(["module"] + tags))

    def ckpt_layer_path(self, ckpt_dir, local_layer_idx):
        stem = f"layer_{local_layer_idx + self._local_start:02d}"
        rank_repr = self._topo.get_rank_repr(rank=self.global_rank)
        if rank_repr:
            stem += f"-{rank_repr}"
        return os.path.join(ckpt_dir, stem + "-model_states.pt")

    def ckpt_layer_path_list(self, ckpt_dir, local_layer_idx):
        return sorted(glob.glob(os.path.join(ckpt_dir, f"layer_{local_layer_idx + self._local_start:02d}-*model_states.pt")))

    def save_state_dict(self, save_dir, checkpoint_engine, exclude_frozen_params=False):
        dp_rank, dp_size = self._grid.data_parallel_id, self._grid.data_parallel_size
        n = len(self.forward_funcs)
        if self.checkpoint_parallel_write_pipeline:
            bounds = ds_utils.partition_uniform(n, dp_size)  # spread the layer files over the data-parallel replicas
            todo = range(bounds[dp_rank], bounds[dp_rank + 1])
        elif dp_rank == 0:
            todo = range(n)
        else:
            return
        checkpoint_engine.makedirs(save_dir, exist_ok=True)
        for k in todo:
            layer = self.forward_funcs[k]
            if not hasattr(layer, "state_dict"):
                continue
            sd = layer.state_dict()
            # clone: torch.save would otherwise serialise the whole flat storage a parameter view lives in
            checkpoint_engine.save(type(sd)((name, t.clone()) for name, t in sd.items()), self.ckpt_layer_path(save_dir, k))

    def load_state_dir(self, load_dir, checkpoint_engine, strict=True):
        mp_rank = self._grid.get_slice_parallel_rank()
        for k, layer in enumerate(self.forward_funcs):
            if not hasattr(layer, "load_state_dict"):
                continue
            files = self.ckpt_layer_path_list(load_dir, k)
            assert len(files) >= 1, f"no checkpoint files for layer {k + self._local_start} in {load_dir}"
            layer.load_state_dict(checkpoint_engine.load(files[min(mp_rank, len(files) - 1)], map_location="cpu"), strict=strict)
        self._synchronize_tied_weights()
