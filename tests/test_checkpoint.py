"""Checkpoint save/load round trip (file layout of the reference) and zero_to_fp32 consolidation, world 2 over gloo."""
import copy
import os
import shutil

import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _ckpt_body(stage, tmp):
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    rank = dist.get_rank()
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0},
           "scheduler": {"type": "WarmupLR", "params": {"warmup_num_steps": 10}}}
    torch.manual_seed(1)
    model = SimpleModel()
    engine, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    batches = random_batches(6, 4, 32, seed=10 + rank)
    for x, y in batches[:3]:
        engine.backward(engine(x, y))
        engine.step()
    engine.save_checkpoint(tmp, client_state={"hello": 7})
    tag = "global_step3"
    files = os.listdir(os.path.join(tmp, tag))
    assert f"zero_pp_rank_{rank}_mp_rank_00_optim_states.pt" in files
    assert ("zero_pp_rank_0_mp_rank_00_model_states.pt" if stage == 3 else "mp_rank_00_model_states.pt") in files
    assert open(os.path.join(tmp, "latest")).read() == tag
    assert os.path.exists(os.path.join(tmp, "zero_to_fp32.py"))
    want = {n: safe_get_full_fp32_param(p).clone() for n, p in model.named_parameters()}
    want_m = {n: safe_get_full_optimizer_state(p, "exp_avg").clone() for n, p in model.named_parameters()}
    # continue training on the original engine
    for x, y in batches[3:]:
        engine.backward(engine(x, y))
        engine.step()
    after = {n: safe_get_full_fp32_param(p).clone() for n, p in model.named_parameters()}
    # fresh engine, load, verify state then verify the continued trajectory is identical
    torch.manual_seed(99)
    model2 = SimpleModel()
    engine2, _, _, _ = deepspeed.initialize(model=model2, config=cfg)
    path, client = engine2.load_checkpoint(tmp)
    assert path is not None and client["hello"] == 7
    assert engine2.global_steps == 3
    for n, p in model2.named_parameters():
        assert torch.allclose(safe_get_full_fp32_param(p), want[n], atol=1e-6), n
        assert torch.allclose(safe_get_full_optimizer_state(p, "exp_avg"), want_m[n], atol=1e-6), n
    for x, y in batches[3:]:
        engine2.backward(engine2(x, y))
        engine2.step()
    for n, p in model2.named_parameters():
        assert torch.allclose(safe_get_full_fp32_param(p), after[n], atol=1e-5), n
    dist.barrier()
    if rank == 0:
        from deepspeedsyclsupport_b200.utils.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
        sd = get_fp32_state_dict_from_zero_checkpoint(tmp)
        for n in want:
            assert torch.allclose(sd[n], want[n].cpu(), atol=1e-6), n


@pytest.mark.parametrize("stage", [1, 3])
def test_checkpoint_roundtrip_two_ranks(stage, tmp_path_factory):
    tmp = str(tmp_path_factory.mktemp(f"ckpt{stage}"))
    run_distributed(_ckpt_body, world_size=2, args=(stage, tmp))


def _universal_save_body(tmp):
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    rank = dist.get_rank()
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}}, "zero_optimization": {"stage": 2}}
    torch.manual_seed(1)
    model = SimpleModel()
    engine, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    for x, y in random_batches(3, 4, 32, seed=10 + rank):
        engine.backward(engine(x, y))
        engine.step()
    engine.save_checkpoint(tmp)
    if rank == 0:
        ref = {n: (safe_get_full_fp32_param(p).clone(), safe_get_full_optimizer_state(p, "exp_avg_sq").clone()) for n, p in model.named_parameters()}
        torch.save(ref, os.path.join(tmp, "expected.pt"))
    else:
        for n, p in model.named_parameters():
            safe_get_full_fp32_param(p), safe_get_full_optimizer_state(p, "exp_avg_sq")  # collective on every rank


def _universal_load_body(tmp, stage):
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}, "checkpoint": {"load_universal": True}}
    torch.manual_seed(5)
    model = SimpleModel()
    engine, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    path, _ = engine.load_checkpoint(tmp)
    assert path is not None and engine.global_steps == 3
    ref = torch.load(os.path.join(tmp, "expected.pt"), weights_only=False)
    for n, p in model.named_parameters():
        w, v = safe_get_full_fp32_param(p), safe_get_full_optimizer_state(p, "exp_avg_sq")
        assert torch.allclose(w, ref[n][0], atol=1e-6), n
        assert torch.allclose(v, ref[n][1], atol=1e-6), n
    assert engine.optimizer._step_count == 3
    x, y = random_batches(1, 4, 32, seed=3)[0]
    engine.backward(engine(x, y))
    engine.step()


def test_universal_checkpoint_reshapes_world_and_stage(tmp_path):
    """save at dp=2 / stage 2 -> ds_to_universal -> load at dp=1 / stage 3 and at dp=2 / stage 1."""
    from deepspeedsyclsupport_b200.checkpoint.ds_to_universal import main as to_universal
    tmp = str(tmp_path)
    run_distributed(_universal_save_body, world_size=2, args=(tmp, ))
    to_universal(["--input_folder", os.path.join(tmp, "global_step3"), "--output_folder", os.path.join(tmp, "global_step3_universal")])
    assert open(os.path.join(tmp, "latest_universal")).read() == "global_step3_universal"
    assert os.path.isfile(os.path.join(tmp, "global_step3_universal", "zero", "head.weight", "exp_avg.pt"))
    run_distributed(_universal_load_body, world_size=1, args=(tmp, 3))
    run_distributed(_universal_load_body, world_size=2, args=(tmp, 1))


def test_reshape_maps():
    from deepspeedsyclsupport_b200.checkpoint import model_3d_desc, reshape_meg_2d_parallel
    m = reshape_meg_2d_parallel(old_pp_degree=2, old_tp_degree=4, new_pp_degree=1, new_tp_degree=2)
    assert m.get_data(0, 0) == [0, 1, 4, 5] and m.get_data(0, 1) == [2, 3, 6, 7]
    maps = model_3d_desc(pp_degree=2, tp_degree=2, dp_degree=4).reshape(model_3d_desc(pp_degree=2, tp_degree=1, dp_degree=2))
    assert len(maps) == 2 and maps[0].get_data(0, 0) == [0, 1, 4, 5] and maps[1].get_data(1, 0) == [10, 11, 14, 15]
    ok, err = model_3d_desc(1, 1, 1).can_reshape(model_3d_desc(2, 1, 1))
    assert not ok and "Expansion" in err[0]


def test_zero_to_fp32_reads_reference_layout_checkpoints(tmp_path):
    """Checkpoints written by the reference implementation carry no `ds_unit_layout`: stage 1/2 = one slice of every param
    group's padded flat buffer per rank, stage 3 = every parameter partitioned by itself (ceil(numel / world) per rank), both
    described by `param_shapes` in the model-states file (deepspeed/utils/zero_to_fp32.py:140-420)."""
    import math
    import os
    from collections import OrderedDict
    from deepspeedsyclsupport_b200.utils.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
    torch.manual_seed(0)
    params = OrderedDict(a=torch.randn(5, 3), b=torch.randn(7), c=torch.randn(2, 2, 2))
    world = 2
    for stage in (2, 3):
        root = tmp_path / f"stage{stage}"
        os.makedirs(root / "global_step1")
        (root / "latest").write_text("global_step1")
        shapes = [OrderedDict((k, v.shape) for k, v in params.items())]
        if stage == 2:
            flat = torch.cat([v.flatten() for v in params.values()])
            flat = torch.cat([flat, flat.new_zeros((-flat.numel()) % (2 * world))])
            per = [[part] for part in flat.chunk(world)]
            model_file = "mp_rank_00_model_states.pt"
        else:
            pieces = [[] for _ in range(world)]
            for v in params.values():
                n = math.ceil(v.numel() / world)
                f = torch.cat([v.flatten(), v.new_zeros(n * world - v.numel())])
                for r in range(world):
                    pieces[r].append(f[r * n:(r + 1) * n])
            per = [[torch.cat(p)] for p in pieces]
            model_file = "zero_pp_rank_0_mp_rank_00_model_states.pt"
        torch.save({"param_shapes": shapes, "buffer_names": [], "shared_params": {}, "module": {}}, root / "global_step1" / model_file)
        key = "fp32_flat_groups" if stage == 3 else "single_partition_of_fp32_groups"
        for r in range(world):
            torch.save({"optimizer_state_dict": {"zero_stage": stage, "partition_count": world, key: per[r]}},
                       root / "global_step1" / f"zero_pp_rank_{r}_mp_rank_00_optim_states.pt")
        sd = get_fp32_state_dict_from_zero_checkpoint(str(root))
        assert all(torch.equal(sd[k], v) for k, v in params.items()), stage
