"""Loading a universal checkpoint (see ds_to_universal.py for the layout) into a live engine at ANY data-parallel world
size / ZeRO stage: every rank reads, per parameter, only the window of the per-parameter files that overlaps its own
slice of the unit the parameter lives in.  Parity: deepspeed/checkpoint/universal_checkpoint.py:12-98 (load_hp_checkpoint_state)
plus the engine-side glue of stage_1_and_2.py:2194-2225 / bf16_optimizer.py:398-430."""
import glob
import os

import torch

from ..utils.logging import log_dist, logger
from .constants import PARAM


def load_hp_checkpoint_state(folder, key="fp32"):
    """The full-shape tensor stored for one parameter (`key` = fp32 | exp_avg | exp_avg_sq | ...)."""
    path = os.path.join(folder, f"{key}.pt")
    obj = torch.load(path, map_location="cpu", weights_only=False)
    return obj[PARAM] if isinstance(obj, dict) else obj


def _fill_rank_slice(opt, unit, name, offset, full_flat, setter):
    """Copy the part of parameter `name` (flat, offset `offset` inside `unit`) that falls into this rank's slice."""
    S = unit.shard_numel
    lo_unit, hi_unit = opt.shard_rank * S, (opt.shard_rank + 1) * S  # my window in unit coordinates
    a, b = max(offset, lo_unit), min(offset + full_flat.numel(), hi_unit)
    if a >= b:
        return
    setter(unit.shard_off + (a - lo_unit), unit.shard_off + (b - lo_unit), full_flat[a - offset:b - offset])


def load_universal_into_optimizer(opt, zero_dir, load_optimizer_states=True):
    """opt: ZeroShardedOptimizer.  Fills master weights, low-precision shard and optimizer moments from <zero_dir>."""
    meta_path = os.path.join(zero_dir, "optimizer_state.pt")
    meta = torch.load(meta_path, map_location="cpu", weights_only=False) if os.path.isfile(meta_path) else {}
    missing = []
    for unit in opt.units:
        for p, offset in zip(unit.params, unit.offsets):
            name = getattr(p, "ds_name", None)
            folder = os.path.join(zero_dir, str(name))
            if name is None or not os.path.isdir(folder):
                missing.append(name)
                continue
            flat = load_hp_checkpoint_state(folder, "fp32").reshape(-1).float()

            def set_master(lo, hi, v):
                for r in opt.regions:
                    x, y = max(lo, r.lo), min(hi, r.hi)
                    if x < y:
                        r.master[x - r.lo:y - r.lo].copy_(v[x - lo:y - lo])
                opt.lp_shard[lo:hi].copy_(v.to(opt.lp_shard.device, opt.lp_shard.dtype))

            _fill_rank_slice(opt, unit, name, offset, flat, set_master)
            if not load_optimizer_states:
                continue
            for state_file in sorted(glob.glob(os.path.join(folder, "*.pt"))):
                key = os.path.basename(state_file)[:-3]
                if key == "fp32":
                    continue
                sflat = load_hp_checkpoint_state(folder, key).reshape(-1).float()

                def set_state(lo, hi, v, key=key):
                    for r in opt.regions:
                        opt._ensure_state(r, opt._opt_kind)
                        x, y = max(lo, r.lo), min(hi, r.hi)
                        if x < y and key in r.state:
                            r.state[key][x - r.lo:y - r.lo].copy_(v[x - lo:y - lo])

                _fill_rank_slice(opt, unit, name, offset, sflat, set_state)
    if missing:
        logger.warning(f"universal checkpoint has no entry for {len(missing)} parameters (first: {missing[:3]})")
    if load_optimizer_states and meta:
        opt._step_count = meta.get("step_count", opt._step_count)
        if meta.get("loss_scaler") is not None:
            opt.loss_scaler = meta["loss_scaler"]
        for g, saved in zip(opt.param_groups, meta.get("param_groups", [])):
            for k, v in saved.items():
                if k != "params":
                    g[k] = v
    opt._post_load_sync()
    return meta


def load_universal_into_engine(engine, load_dir, tag, load_optimizer_states=True, load_lr_scheduler_states=True):
    """engine.load_checkpoint() path for `"checkpoint": {"load_universal": true}`."""
    ckpt_dir = os.path.join(load_dir, str(tag))
    zero_dir = os.path.join(ckpt_dir, "zero")
    assert os.path.isdir(zero_dir), f"{zero_dir} is not a universal checkpoint (run ds_to_universal first)"
    opts = engine.optimizer._all() if hasattr(engine.optimizer, "_all") else [engine.optimizer]
    for opt in opts:
        load_universal_into_optimizer(opt, zero_dir, load_optimizer_states=load_optimizer_states)
    client_state = {}
    ms_path = os.path.join(ckpt_dir, "mp_rank_00_model_states.pt")
    if os.path.isfile(ms_path):
        ms = torch.load(ms_path, map_location="cpu", weights_only=False)
        for n, b in engine.module.named_buffers():
            if n in ms.get("module", {}) and ms["module"][n].numel() == b.numel():
                b.data.copy_(ms["module"][n].to(b.device))
        if load_lr_scheduler_states and engine.lr_scheduler is not None and ms.get("lr_scheduler") is not None:
            engine.lr_scheduler.load_state_dict(ms["lr_scheduler"])
        engine.global_steps = ms.get("global_steps", 0)
        engine.global_samples = ms.get("global_samples", engine.global_steps * engine.train_batch_size())
        engine.skipped_steps = ms.get("skipped_steps", 0)
        known = {"module", "buffer_names", "optimizer", "param_shapes", "frozen_param_shapes", "frozen_param_fragments", "shared_params", "lr_scheduler",
                 "data_sampler", "random_ltd", "sparse_tensor_module_names", "skipped_steps", "global_steps", "global_samples", "dp_world_size",
                 "mp_world_size", "ds_config", "ds_version", "ds_lp_shard"}
        client_state = {k: v for k, v in ms.items() if k not in known}
    log_dist(f"loaded universal checkpoint {ckpt_dir}", ranks=[0])
    return ckpt_dir, client_state
