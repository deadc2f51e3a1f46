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


def test_moq_quantizer_lowers_bits_over_periods():
    from deepspeedsyclsupport_b200.runtime.quantize import Quantizer
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.randn(16, 32))
    w.start_bits, w.target_bits, w.q_period = 8, 4, 2
    q = Quantizer(q_groups=4, q_mixed_fp16=False, q_type="symmetric", q_rounding="nearest")
    levels = []
    for step in range(40):
        w.data = torch.randn(16, 32)
        q.quantize([[w]], overflow=False, eigenvalue_enabled=False)
        levels.append((w.start_bits, w.data.reshape(4, -1)[0].unique().numel()))
    bits = [b for b, _ in levels]
    assert bits[0] == 8 and bits[-1] == 4 and sorted(set(bits), reverse=True) == [8, 7, 6, 5, 4]
    assert bits.index(7) == 1 and bits.index(6) == 3 and bits.index(5) == 7 and bits.index(4) == 15  # the period doubles at every drop
    assert all(n <= 2**b for b, n in levels)
    # an eigenvalue stretches the next period: factor 1 + floor(4 * 0.5) = 3
    v = torch.nn.Parameter(torch.randn(8, 8))
    v.start_bits, v.target_bits, v.q_period = 8, 6, 1
    q2 = Quantizer(q_groups=1)
    q2.quantize([[v]], False, True, {id(v): (0.5, 0)})
    assert v.start_bits == 7 and v.q_period == 2 * 3
    # overflow skips the step (unless eigenvalues drive the schedule); ternary / binary forms
    q2.quantize([[v]], True, False)
    assert q2.qsteps == 1
    t = Quantizer(q_groups=2).quantize_tenary(torch.randn(4, 16))
    assert all(g.unique().numel() <= 3 for g in t.reshape(2, -1))
    b = Quantizer(q_groups=2).quantize_binary(torch.randn(4, 16))
    assert all(g.abs().unique().numel() == 1 for g in b.reshape(2, -1))
    # mixed-precision blending moves from unquantised to quantised at q_change_ratio per step
    m = torch.nn.Parameter(torch.randn(4, 8))
    m.start_bits, m.target_bits, m.q_period = 3, 3, 100
    qm = Quantizer(q_groups=1, q_mixed_fp16=True, q_change_ratio=0.5)
    orig = m.data.clone()
    qm.quantize([[m]], False, False)
    half = m.data.clone()
    full = Quantizer(q_groups=1).quantize_highbit(orig, 3)
    assert torch.allclose(half, 0.5 * orig + 0.5 * full, atol=1e-6)
