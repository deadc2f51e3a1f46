"""BASELINE.json config #1: GPT-2 (small architecture, reduced width for test time), ZeRO-1, gloo, world_size=2 --
loss parity with plain PyTorch data parallelism (manual gradient averaging = what DDP computes)."""
import copy

import torch

from common import run_distributed


def _body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models import GPT2LMHeadModel, gpt2_config
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    cfg = gpt2_config("small", n_embd=96, n_head=4, n_layer=3, vocab_size=211, n_positions=32)
    model = GPT2LMHeadModel(cfg)
    ref = copy.deepcopy(model)
    engine, _, _, _ = deepspeed.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.01}},
        "zero_optimization": {"stage": 1}, "gradient_clipping": 1.0})
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-3, weight_decay=0.01)
    gens = [torch.Generator().manual_seed(50 + r) for r in range(world)]
    for step in range(5):
        batches = [(torch.randint(0, 211, (2, 32), generator=g), torch.randint(0, 211, (2, 32), generator=g)) for g in gens]
        ids, labels = batches[rank]
        loss = engine(ids, labels)
        engine.backward(loss)
        engine.step()
        ropt.zero_grad()
        ref_losses = []
        for r in range(world):
            l = ref(*batches[r])
            (l / world).backward()
            ref_losses.append(l.item())
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        ropt.step()
        assert abs(loss.item() - ref_losses[rank]) < 2e-4, (step, loss.item(), ref_losses[rank])


def test_gpt2_zero1_gloo_world2_loss_parity():
    run_distributed(_body, world_size=2)
