"""BF16_Optimizer / FP16_Optimizer class-name wrappers and PLD / eigenvalue helpers (CPU)."""
import torch

from simple_model import SimpleModel, random_batches


def _train(opt, model, dtype, n=6):
    losses = []
    for x, y in random_batches(n, 8, 16, seed=1, dtype=dtype):
        loss = model(x, y)
        opt.backward(loss)
        opt.end_micro_step(True)
        opt.step()
        losses.append(loss.item())
    return losses


def test_bf16_optimizer_trains_and_keeps_fp32_master():
    from deepspeedsyclsupport_b200.ops.adam import FusedAdam
    from deepspeedsyclsupport_b200.runtime.bf16_optimizer import BF16_Optimizer
    torch.manual_seed(0)
    model = SimpleModel(16, 2).bfloat16()
    opt = BF16_Optimizer(FusedAdam(model.parameters(), lr=1e-2), module=model, clip_grad=1.0)
    losses = _train(opt, model, torch.bfloat16)
    assert losses[-1] < losses[0]
    p = next(model.parameters())
    assert p.dtype == torch.bfloat16 and opt.get_full_fp32_param(p).dtype == torch.float32
    assert torch.allclose(opt.get_full_fp32_param(p).bfloat16().float(), p.float(), atol=1e-2)


def test_fp16_optimizer_dynamic_scale_skips_overflow():
    from deepspeedsyclsupport_b200.ops.adam import FusedAdam
    from deepspeedsyclsupport_b200.runtime.fp16.fused_optimizer import FP16_Optimizer
    torch.manual_seed(0)
    model = SimpleModel(16, 2).half()
    opt = FP16_Optimizer(FusedAdam(model.parameters(), lr=1e-3), module=model, dynamic_loss_scale=True,
                         dynamic_loss_args={"init_scale": 2**8, "scale_window": 2, "min_scale": 1, "delayed_shift": 1})
    before = [p.detach().clone() for p in model.parameters()]
    x, y = random_batches(1, 8, 16, seed=1, dtype=torch.float16)[0]
    loss = model(x, y)
    opt.backward(loss)
    opt.end_micro_step(True)
    opt.grad_shard[0] = float("inf")
    opt.step()
    assert opt.overflow and opt.cur_scale == 2**7
    assert all(torch.equal(a, b) for a, b in zip(before, model.parameters()))
    losses = _train(opt, model, torch.float16)
    assert not opt.overflow and losses[-1] < losses[0]


def test_pld_and_eigenvalue():
    from deepspeedsyclsupport_b200.runtime.eigenvalue import Eigenvalue
    from deepspeedsyclsupport_b200.runtime.progressive_layer_drop import ProgressiveLayerDrop
    pld = ProgressiveLayerDrop(theta=0.5, gamma=0.01)
    pld.update_state(0)
    assert pld.get_theta() == 1.0
    pld.update_state(10000)
    assert abs(pld.get_theta() - 0.5) < 1e-6 and pld.get_state()["progressive_layer_drop"]
    torch.manual_seed(0)
    model = SimpleModel(8, 2).double()
    x, y = random_batches(1, 4, 8, seed=3, dtype=torch.float64)[0]
    x = torch.tanh(x)
    h = x
    for l in model.linears:
        h = torch.tanh(l(h))
    loss = (h**2).sum()
    loss.backward(create_graph=True)
    ev = Eigenvalue(layer_name="linears", layer_num=2, max_iter=50, tol=1e-3).compute_eigenvalue(model)
    assert len(ev) == 4 and all(0 <= v[0] <= 1.0 for v in ev.values()) and max(v[0] for v in ev.values()) == 1.0
