#!/usr/bin/env python
"""Consolidate a ZeRO checkpoint (any stage) into a single fp32 state dict -- standalone (needs only torch); a copy is
placed next to every checkpoint.  Usage:   python zero_to_fp32.py <checkpoint_dir> <output_file> [-t TAG]

Parity: deepspeed/utils/zero_to_fp32.py:33-592 (same three public functions and CLI).  Layout difference: each rank
file `[bf16_]zero_pp_rank_R_mp_rank_00_optim_states.pt` holds ONE flat fp32 shard (`fp32_flat_groups` for stage 3,
`single_partition_of_fp32_groups` otherwise) plus `ds_unit_layout`, which lists the units (contiguous groups of
parameters) with, per unit, its padded size and every parameter's (name, shape, offset).  Rank r owns slice r of
every unit; concatenating the rank slices of a unit gives the unit's flat buffer.
"""
import argparse
import glob
import math
import os
import re
from collections import OrderedDict

import torch


def _natural_keys(text):
    return [int(c) if c.isdigit() else c for c in re.split(r"(\d+)", text)]


def get_optim_files(checkpoint_dir):
    files = sorted(glob.glob(os.path.join(checkpoint_dir, "*_optim_states.pt")), key=_natural_keys)
    files = [f for f in files if "zero_pp_rank" in os.path.basename(f)]
    if len(files) == 0:
        raise FileNotFoundError(f"can't find '*zero_pp_rank*_optim_states.pt' files in directory '{checkpoint_dir}'")
    return files


def get_model_state_file(checkpoint_dir, zero_stage):
    if zero_stage == 3:
        file = os.path.join(checkpoint_dir, "zero_pp_rank_0_mp_rank_00_model_states.pt")
        if os.path.exists(file):
            return file
    file = os.path.join(checkpoint_dir, "mp_rank_00_model_states.pt")
    if not os.path.exists(file):
        raise FileNotFoundError(f"can't find model states file at '{file}'")
    return file


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False)


def _merge_optimizer_sd(osd):
    """Flatten a (possibly composite MoE) optimizer state dict into a list of plain ZeRO state dicts."""
    if "dense" in osd:
        return [osd["dense"]] + list(osd.get("expert", {}).values())
    return [osd]


def _get_fp32_state_dict_from_zero_checkpoint(ds_checkpoint_dir):
    optim_files = get_optim_files(ds_checkpoint_dir)
    per_rank = [_load(f)["optimizer_state_dict"] for f in optim_files]
    first = _merge_optimizer_sd(per_rank[0])[0]
    zero_stage = first["zero_stage"]
    world = first["partition_count"]
    if world != len(optim_files):
        raise ValueError(f"Expected {world} of '*_optim_states.pt' under '{ds_checkpoint_dir}' but found {len(optim_files)} files. "
                         "Possibly due to an overwrite of an old checkpoint, or a checkpoint didn't get saved by one or more processes.")
    print(f"Detected checkpoint of type zero stage {zero_stage}, world_size: {world}")
    state_dict = OrderedDict()
    if "ds_unit_layout" not in first:
        # written by the reference implementation: per-param-group flat partitions described by `param_shapes`
        state_dict = _from_reference_layout(ds_checkpoint_dir, per_rank, zero_stage, world)
        print(f"Reconstructed fp32 state dict with {len(state_dict)} params {sum(v.numel() for v in state_dict.values())} elements")
        return state_dict
    n_sub = len(_merge_optimizer_sd(per_rank[0]))
    for sub in range(n_sub):
        sds = [_merge_optimizer_sd(r)[sub] for r in per_rank]
        layout = sds[0]["ds_unit_layout"]
        key = "fp32_flat_groups" if "fp32_flat_groups" in sds[0] else "single_partition_of_fp32_groups"
        shards = [sd[key][0].float() for sd in sds]
        for u in layout["units"]:
            S, off = u["shard_numel"], u["shard_off"]
            flat = torch.cat([shards[r][off:off + S] for r in range(layout["shard_world"])])
            for name, shape, poff in u["params"]:
                n = int(math.prod(shape)) if len(shape) else 1
                state_dict[name] = flat[poff:poff + n].view(*shape).clone()
    # buffers + shared (tied) parameters come from the model states file
    try:
        ms = _load(get_model_state_file(ds_checkpoint_dir, zero_stage))
        for bn in ms.get("buffer_names", []) or []:
            if bn in ms["module"]:
                state_dict[bn] = ms["module"][bn].float() if ms["module"][bn].is_floating_point() else ms["module"][bn]
        for dst, src in (ms.get("shared_params") or {}).items():
            if src in state_dict:
                state_dict[dst] = state_dict[src]
    except FileNotFoundError:
        pass
    print(f"Reconstructed fp32 state dict with {len(state_dict)} params {sum(v.numel() for v in state_dict.values())} elements")
    return state_dict


def _from_reference_layout(ds_checkpoint_dir, per_rank, zero_stage, world):
    """Checkpoints of the reference DeepSpeed (no `ds_unit_layout`): the model-states file lists, per optimizer param group, the
    parameter names and shapes in flattening order (`param_shapes`).
      stage 1/2  every rank holds one contiguous slice of each group's flat fp32 buffer (the buffer is padded at its end so it
                 divides evenly): concatenate the slices, then cut the parameters off one after the other;
      stage 3    every parameter is partitioned by itself: each rank's flat buffer holds, parameter after parameter,
                 ceil(numel / world) elements of it (the last rank's piece is padded)."""
    ms = _load(get_model_state_file(ds_checkpoint_dir, zero_stage))
    param_shapes = ms["param_shapes"]
    if isinstance(param_shapes, dict):
        param_shapes = [param_shapes]
    key = "fp32_flat_groups" if zero_stage == 3 else "single_partition_of_fp32_groups"
    groups_per_rank = [[t.float().reshape(-1) for t in r[key]] for r in per_rank]
    out = OrderedDict()
    if zero_stage == 3:
        # the reference concatenates all groups of a rank before walking the parameters
        flat_per_rank = [torch.cat(g) for g in groups_per_rank]
        offset = 0
        for shapes in param_shapes:
            for name, shape in shapes.items():
                numel = int(math.prod(shape)) if len(shape) else 1
                piece = math.ceil(numel / world)
                full = torch.cat([flat_per_rank[r][offset:offset + piece] for r in range(world)])
                out[name] = full[:numel].view(*shape).clone()
                offset += piece
    else:
        for g, shapes in enumerate(param_shapes):
            full = torch.cat([groups_per_rank[r][g] for r in range(world)])
            offset = 0
            for name, shape in shapes.items():
                numel = int(math.prod(shape)) if len(shape) else 1
                out[name] = full[offset:offset + numel].view(*shape).clone()
                offset += numel
    for name, value in (ms.get("frozen_param_fragments") or {}).items():
        out.setdefault(name, value.float() if value.is_floating_point() else value)
    for bn in ms.get("buffer_names", []) or []:
        if bn in ms.get("module", {}):
            out[bn] = ms["module"][bn]
    for dst, src in (ms.get("shared_params") or {}).items():
        if src in out:
            out[dst] = out[src]
    return out


def get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag=None, exclude_frozen_parameters=False):
    """Convert ZeRO 1/2/3 checkpoint into a single fp32 consolidated state_dict (on CPU)."""
    if tag is None:
        latest_path = os.path.join(checkpoint_dir, "latest")
        if os.path.isfile(latest_path):
            with open(latest_path, "r") as fd:
                tag = fd.read().strip()
        else:
            raise ValueError(f"Unable to find 'latest' file at {latest_path}")
    ds_checkpoint_dir = os.path.join(checkpoint_dir, tag)
    if not os.path.isdir(ds_checkpoint_dir):
        raise FileNotFoundError(f"Directory '{ds_checkpoint_dir}' doesn't exist")
    return _get_fp32_state_dict_from_zero_checkpoint(ds_checkpoint_dir)


def convert_zero_checkpoint_to_fp32_state_dict(checkpoint_dir, output_file, tag=None, exclude_frozen_parameters=False):
    state_dict = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters)
    print(f"Saving fp32 state dict to {output_file}")
    torch.save(state_dict, output_file)


def load_state_dict_from_zero_checkpoint(model, checkpoint_dir, tag=None):
    """Load the consolidated fp32 weights into `model` (on CPU) and return it."""
    print("Extracting fp32 weights")
    state_dict = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag)
    print("Overwriting model with fp32 weights")
    model = model.cpu()
    model.load_state_dict(state_dict, strict=False)
    return model


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("checkpoint_dir", type=str, help="path to the desired checkpoint folder, e.g., path/checkpoint-12")
    parser.add_argument("output_file", type=str, help="path to the pytorch fp32 state_dict output file (e.g. path/checkpoint-12/pytorch_model.bin)")
    parser.add_argument("-t", "--tag", type=str, default=None, help="checkpoint tag used as a unique identifier for checkpoint. e.g., global_step1")
    parser.add_argument("--exclude_frozen_parameters", action="store_true", help="exclude frozen parameters")
    parser.add_argument("-d", "--debug", action="store_true", help="enable debug")
    args = parser.parse_args()
    convert_zero_checkpoint_to_fp32_state_dict(args.checkpoint_dir, args.output_file, tag=args.tag,
                                               exclude_frozen_parameters=args.exclude_frozen_parameters)
