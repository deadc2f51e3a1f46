"""ZeRO 0/1/2/3 parity against plain PyTorch (world 1 in-process, world 2 over gloo)."""
import copy

import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _train(stage, world_checks=True, gas=2, steps=6, clip=1.0, opt_type="AdamW", extra_zero=None):
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(1234)
    model = SimpleModel()
    ref = copy.deepcopy(model)
    zero = {"stage": stage}
    zero.update(extra_zero or {})
    cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": gas,
           "optimizer": {"type": opt_type, "params": {"lr": 1e-2, "weight_decay": 0.01}},
           "zero_optimization": zero, "gradient_clipping": clip}
    eng, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.01)
    all_batches = [random_batches(steps, 4, 32, seed=100 + r) for r in range(world)]
    for it in range(steps):
        x, y = all_batches[rank][it]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        # reference: average of all ranks' micro-batch gradients
        for r in range(world):
            xr, yr = all_batches[r][it]
            (ref(xr, yr) / (gas * world)).backward()
        if (it + 1) % gas == 0:
            if clip > 0:
                torch.nn.utils.clip_grad_norm_(ref.parameters(), clip)
            ropt.step()
            ropt.zero_grad()
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        full = safe_get_full_fp32_param(p)
        assert torch.allclose(full.cpu(), q.detach(), atol=2e-5, rtol=1e-4), f"stage {stage} param {n} differs: {(full.cpu() - q).abs().max()}"


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
def test_zero_single_process(stage):
    run_distributed(_train, world_size=1, args=(stage, ))


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
def test_zero_two_ranks(stage):
    run_distributed(_train, world_size=2, args=(stage, ))


def test_zero3_no_persistence_two_ranks():
    run_distributed(_train, world_size=2, args=(3, True, 1, 4, 0.0, "AdamW", {"stage3_param_persistence_threshold": 0}))
