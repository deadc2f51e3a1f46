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
