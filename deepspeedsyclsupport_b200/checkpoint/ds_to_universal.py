#!/usr/bin/env python
"""Convert a ZeRO checkpoint (stage 0-3, any data-parallel world size) into the universal layout, which can be loaded at
a different world size / stage with `"checkpoint": {"load_universal": true}`:

    <output>/zero/<param name>/fp32.pt          {"param": fp32 tensor of the parameter's full shape}
    <output>/zero/<param name>/exp_avg.pt        one file per optimizer state tensor
    <output>/zero/<param name>/exp_avg_sq.pt
    <output>/zero/optimizer_state.pt             step count, param-group hyper-parameters, loss scaler
    <output>/mp_rank_00_model_states.pt          copy of the module states file (buffers, lr scheduler, counters)
and <folder>/latest_universal names the tag.

How: each rank file holds ONE flat shard per state plus `ds_unit_layout`; for every unit the rank slices are joined in
rank order, then every parameter is a (offset, shape) window of that unit -- the same walk as zero_to_fp32, done for
fp32 weights AND optimizer moments, streamed one unit at a time so host memory stays at one unit x (1 + #states).
CLI parity: deepspeed/checkpoint/ds_to_universal.py (--input_folder, --output_folder, --num_extract_workers,
--num_merge_workers, --keep_temp_folder)."""
import argparse
import glob
import math
import os
import re
import shutil

import torch

from .constants import (DS_UNIT_LAYOUT, FP32_FLAT_GROUPS, OPTIMIZER_STATE_DICT, PARAM, SINGLE_PARTITION_OF_FP32_GROUPS,
                        UNIVERSAL_CHECKPOINT_INFO, UNIVERSAL_CHECKPOINT_VERSION_KEY, UNIVERSAL_CHECKPOINT_VERSION_VALUE)


def parse_arguments(args=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--input_folder", type=str, required=True, help="Input DeepSpeed Checkpoint folder (the tag directory)")
    parser.add_argument("--output_folder", type=str, required=True, help="Output DeepSpeed checkpoint folder")
    parser.add_argument("--num_extract_workers", default=4, type=int, help="kept for CLI compatibility (conversion is streamed per unit)")
    parser.add_argument("--num_merge_workers", default=2, type=int, help="kept for CLI compatibility")
    parser.add_argument("--keep_temp_folder", action="store_true", help="kept for CLI compatibility (no temp folder is produced)")
    parser.add_argument("--no_strict", dest="strict", action="store_false", help="Do not perform validity checks on converted checkpoint.")
    return parser.parse_args(args)


def _natural(text):
    return [int(c) if c.isdigit() else c for c in re.split(r"(\d+)", text)]


def _optim_files(folder):
    files = [f for f in glob.glob(os.path.join(folder, "*_optim_states.pt")) if "zero_pp_rank" in os.path.basename(f)]
    if not files:
        raise FileNotFoundError(f"no *zero_pp_rank*_optim_states.pt files under {folder}")
    return sorted(files, key=_natural)


def _sub_state_dicts(osd):
    """A MoE checkpoint nests one ZeRO state dict for the dense parameters and one per expert group."""
    return [osd["dense"]] + list(osd.get("expert", {}).values()) if "dense" in osd else [osd]


def _numel(shape):
    return int(math.prod(shape)) if len(shape) else 1


def convert(input_folder, output_folder, strict=True):
    files = _optim_files(input_folder)
    per_rank = [_sub_state_dicts(torch.load(f, map_location="cpu", weights_only=False)[OPTIMIZER_STATE_DICT]) for f in files]
    world = per_rank[0][0]["partition_count"]
    if strict:
        assert world == len(files), f"checkpoint was written by {world} shard ranks but {len(files)} optimizer files were found"
    zero_dir = os.path.join(output_folder, "zero")
    os.makedirs(zero_dir, exist_ok=True)
    meta = {"step_count": per_rank[0][0].get("step_count", 0), "loss_scaler": per_rank[0][0].get("loss_scaler"),
            "param_groups": per_rank[0][0][OPTIMIZER_STATE_DICT]["param_groups"], "zero_stage": per_rank[0][0]["zero_stage"], "source_world": world,
            UNIVERSAL_CHECKPOINT_INFO: {UNIVERSAL_CHECKPOINT_VERSION_KEY: UNIVERSAL_CHECKPOINT_VERSION_VALUE}}
    n_params = 0
    for sub in range(len(per_rank[0])):
        sds = [r[sub] for r in per_rank]
        layout = sds[0][DS_UNIT_LAYOUT]
        weight_key = FP32_FLAT_GROUPS if FP32_FLAT_GROUPS in sds[0] else SINGLE_PARTITION_OF_FP32_GROUPS
        streams = {"fp32": [sd[weight_key][0] for sd in sds]}
        for state_name in sds[0][OPTIMIZER_STATE_DICT]["state"].get(0, {}):
            streams[state_name] = [sd[OPTIMIZER_STATE_DICT]["state"][0][state_name] for sd in sds]
        shard_world = layout["shard_world"]
        for unit in layout["units"]:
            S, off = unit["shard_numel"], unit["shard_off"]
            joined = {k: torch.cat([v[r][off:off + S].float() for r in range(shard_world)]) for k, v in streams.items()}
            for name, shape, poff in unit["params"]:
                pdir = os.path.join(zero_dir, name)
                os.makedirs(pdir, exist_ok=True)
                n = _numel(shape)
                for k, flat in joined.items():
                    torch.save({PARAM: flat[poff:poff + n].view(*shape).clone()}, os.path.join(pdir, f"{k}.pt"))
                n_params += 1
    torch.save(meta, os.path.join(zero_dir, "optimizer_state.pt"))
    for f in glob.glob(os.path.join(input_folder, "*mp_rank_*_model_states.pt")):
        if "expert" in os.path.basename(f) or "zero_pp_rank" in os.path.basename(f) and not os.path.basename(f).startswith("zero_pp_rank_0_"):
            continue
        shutil.copy(f, os.path.join(output_folder, os.path.basename(f).replace("zero_pp_rank_0_", "")))
    for f in glob.glob(os.path.join(input_folder, "layer_*_model_states.pt")):
        shutil.copy(f, output_folder)
    return n_params


def main(args=None):
    args = parse_arguments(args)
    print(f"Convert DeepSpeed Checkpoint to Universal Checkpoint\nConverting DeepSpeed checkpoint in {args.input_folder} to Universal checkpoint in "
          f"{args.output_folder}")
    n = convert(args.input_folder, args.output_folder, strict=args.strict)
    # <parent>/latest_universal so engine.load_checkpoint(load_universal) finds the converted tag
    parent, tag = os.path.split(os.path.normpath(args.output_folder))
    with open(os.path.join(parent, "latest_universal"), "w") as f:
        f.write(tag)
    print(f"*** Done: {n} parameters written")


if __name__ == "__main__":
    main()
