"""MoE: gating semantics vs the dense one-hot formulation, and expert-parallel training on 2 gloo ranks."""
import pytest
import torch

from common import run_distributed


@pytest.mark.parametrize("k", [1, 2])
def test_gating_dispatch_combine_matches_dense_einsum(k):
    from deepspeedsyclsupport_b200.moe import sharded_moe as M
    torch.manual_seed(0)
    S, E, D = 64, 4, 8
    x = torch.randn(S, D)
    gate = M.TopKGate(D, E, k=k, capacity_factor=1.0, min_capacity=4, use_rts=False)
    l_aux, te, ts, tw, cap, counts = gate(x)
    assert l_aux.item() > 0 and te.shape == (S, k)
    cw, dm = M.dense_dispatch_masks(te, ts, tw, E, cap)
    dispatched_dense = torch.einsum("sec,sm->ecm", dm.float(), x)
    dispatched = M.dispatch_tokens(x, te, ts, E, cap)
    assert torch.allclose(dispatched, dispatched_dense, atol=1e-6)
    eo = torch.randn(E, cap, D)
    combined_dense = torch.einsum("sec,ecm->sm", cw, eo)
    combined = M.combine_tokens(eo, te, ts, tw, S)
    assert torch.allclose(combined, combined_dense, atol=1e-5)
    # capacity respected, slots unique per expert
    for e in range(E):
        slots = ts[te == e]
        assert slots.numel() <= cap and slots.unique().numel() == slots.numel()


class _Expert(torch.nn.Module):

    def __init__(self, d):
        super().__init__()
        self.fc = torch.nn.Linear(d, d)

    def forward(self, x):
        return torch.relu(self.fc(x))


class _MoEModel(torch.nn.Module):

    def __init__(self, d, num_experts, ep_size):
        super().__init__()
        from deepspeedsyclsupport_b200.moe.layer import MoE
        self.inp = torch.nn.Linear(d, d)
        self.moe = MoE(hidden_size=d, expert=_Expert(d), num_experts=num_experts, ep_size=ep_size, k=1, use_rts=False,
                       capacity_factor=2.0, min_capacity=0)
        self.out = torch.nn.Linear(d, 4)

    def forward(self, x, y):
        h = torch.relu(self.inp(x))
        h, l_aux, _ = self.moe(h)
        return torch.nn.functional.cross_entropy(self.out(h), y) + 0.01 * l_aux


def _ep_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.moe.utils import is_moe_param
    torch.manual_seed(0)
    model = _MoEModel(16, num_experts=4, ep_size=2)
    cfg = {"train_micro_batch_size_per_gpu": 8, "optimizer": {"type": "AdamW", "params": {"lr": 5e-3}},
           "zero_optimization": {"stage": 2}, "gradient_clipping": 1.0}
    engine, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    assert engine.has_moe_layers
    assert sum(1 for p in model.parameters() if is_moe_param(p)) == 2 * 2  # 2 local experts x (w, b)
    g = torch.Generator().manual_seed(11 + dist.get_rank())
    x, y = torch.randn(8, 16, generator=g), torch.randint(0, 4, (8, ), generator=g)
    losses = []
    for _ in range(25):
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0], losses
    import tempfile, os
    d = os.environ.get("DSB200_TEST_TMP", "/tmp/dsb200_moe_ckpt")
    engine.save_checkpoint(d, tag="t1")
    dist.barrier()
    assert os.path.exists(os.path.join(d, "t1", f"layer_0_expert_{dist.get_rank() * 2}_mp_rank_00_model_states.pt"))
    assert os.path.exists(os.path.join(d, "latest"))
    # load into a differently initialised engine: dense parameters, this rank's experts and the optimizer state come back
    torch.manual_seed(99)
    model2 = _MoEModel(16, num_experts=4, ep_size=2)
    engine2, _, _, _ = deepspeed.initialize(model=model2, config=cfg)
    before = engine2(x, y).item()
    path, client = engine2.load_checkpoint(d, tag="t1")
    assert path is not None and engine2.global_steps == engine.global_steps
    for (n, p), (_, q) in zip(model.named_parameters(), model2.named_parameters()):
        assert torch.allclose(p.detach().float(), q.detach().float(), atol=1e-6), n
    engine.eval(), engine2.eval()
    with torch.no_grad():
        assert abs(engine2(x, y).item() - engine(x, y).item()) < 1e-5 and abs(before - engine(x, y).item()) > 1e-4
    engine.train(), engine2.train()
    for e in (engine, engine2):  # one more identical step on both: the restored optimizer state matches too
        loss = e(x, y)
        e.backward(loss)
        e.step()
    for (n, p), (_, q) in zip(model.named_parameters(), model2.named_parameters()):
        assert torch.allclose(p.detach().float(), q.detach().float(), atol=1e-5), n


def test_moe_expert_parallel_two_ranks():
    run_distributed(_ep_body, world_size=2)


def test_grouped_experts_match_the_expert_loop_forward_and_backward(monkeypatch):
    """Two-layer MLP experts are detected by probing and run as grouped GEMMs; outputs and every gradient equal the per-expert
    loop.  Experts the probe cannot prove (extra parameters, dropout) keep the loop."""
    from torch import nn
    from deepspeedsyclsupport_b200.moe.experts import Experts

    class MLP(nn.Module):

        def __init__(self, act):
            super().__init__()
            self.up, self.down, self.act = nn.Linear(16, 40), nn.Linear(40, 16), act

        def forward(self, x):
            return self.down(self.act(self.up(x)))

    torch.manual_seed(0)
    for act, name in ((nn.GELU(), "gelu"), (nn.ReLU(), "relu"), (nn.SiLU(), "silu")):
        grouped = Experts(MLP(act), num_local_experts=3)
        assert grouped._grouped_act is not None and grouped._grouped_act[0] == name
        with torch.no_grad():
            for e in grouped.deepspeed_experts:  # deepcopy made the experts identical: make them differ
                for p in e.parameters():
                    p.add_(0.1 * torch.randn_like(p))
        monkeypatch.setenv("DSB200_GROUPED_EXPERTS", "0")
        loop = Experts(MLP(act), num_local_experts=3)
        monkeypatch.delenv("DSB200_GROUPED_EXPERTS")
        assert loop._grouped_act is None
        loop.load_state_dict(grouped.state_dict())
        x = torch.randn(2, 3, 5, 16, requires_grad=True)  # [ep, local experts, capacity, model dim]
        x2 = x.detach().clone().requires_grad_(True)
        w = torch.randn(2, 3, 5, 16)
        (grouped(x) * w).sum().backward()
        (loop(x2) * w).sum().backward()
        assert torch.allclose(grouped(x), loop(x2), atol=1e-5) and torch.allclose(x.grad, x2.grad, atol=1e-5)
        for (n, p), (_, q) in zip(grouped.named_parameters(), loop.named_parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-5), n

    class WithNorm(MLP):

        def __init__(self):
            super().__init__(nn.GELU())
            self.norm = nn.LayerNorm(16)

        def forward(self, x):
            return super().forward(self.norm(x))

    class WithDropout(MLP):

        def __init__(self):
            super().__init__(nn.GELU())
            self.drop = nn.Dropout(0.1)

    class Residual(MLP):

        def forward(self, x):
            return x + super().forward(x)

    for other in (WithNorm(), WithDropout(), Residual(nn.GELU())):
        assert Experts(other, num_local_experts=2)._grouped_act is None
