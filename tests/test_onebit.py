"""Compressed all-reduce numerics + the three communication-compressed optimizers under the engine (gloo, 2 ranks).
Reference tests: tests/unit/runtime/half_precision/onebit/test_onebit.py, tests/onebit/test_nccl_backend.py."""
import numpy as np
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _sim_compressed_allreduce(xs, worker_errors, server_errors):
    """Plain-torch model of the algorithm for `world` simulated ranks (the formula of the upstream unit test)."""
    world = len(xs)
    n = xs[0].numel()
    signs, scales = [], []
    for r in range(world):
        v = xs[r] + worker_errors[r]
        s = torch.linalg.norm(v) / np.sqrt(n)
        sg = torch.where(v >= 0, 1.0, -1.0)
        worker_errors[r] = v - s * sg
        signs.append(sg)
        scales.append(s)
    chunk = n // world
    out = []
    for r in range(world):
        m = sum(signs[w][r * chunk:(r + 1) * chunk] * scales[w] for w in range(world)) / world + server_errors[r]
        ss = torch.linalg.norm(m) / np.sqrt(chunk)
        sg = torch.where(m >= 0, 1.0, -1.0)
        server_errors[r] = m - ss * sg
        out.append(ss * sg)
    return torch.cat(out)


def _allreduce_body():
    import deepspeedsyclsupport_b200.comm as dist
    from deepspeedsyclsupport_b200.runtime.comm.nccl import NcclBackend
    rank, world = dist.get_rank(), dist.get_world_size()
    backend = NcclBackend()
    n = 16 * world * 8
    gen = torch.Generator().manual_seed(0)
    xs = [torch.randn(n, generator=gen) for _ in range(world)]
    we = [torch.zeros(n) for _ in range(world)]
    se = [torch.zeros(n // world) for _ in range(world)]
    my_we, my_se = torch.zeros(n), torch.zeros(n // world)
    for it in range(3):
        expect = _sim_compressed_allreduce([x.clone() for x in xs], we, se)
        got = backend.compressed_allreduce(xs[rank].clone(), my_we, my_se, 0)
        assert torch.allclose(got, expect, atol=1e-6), (it, (got - expect).abs().max())
        assert torch.allclose(my_we, we[rank], atol=1e-6) and torch.allclose(my_se, se[rank], atol=1e-6)
        xs = [x * 0.9 + 0.1 * torch.randn(n, generator=gen) for x in xs]


def test_compressed_allreduce_matches_formula():
    run_distributed(_allreduce_body, world_size=2)


def _train_body(opt_type, opt_params):
    import deepspeedsyclsupport_b200 as deepspeed
    import deepspeedsyclsupport_b200.comm as dist
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=16, nlayers=2)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": opt_type, "params": opt_params}, "gradient_clipping": 1.0}
    engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=cfg)
    rank = dist.get_rank()
    losses = []
    saw_compressed = False
    for i, (x, y) in enumerate(random_batches(12, 4, 16, seed=100 + rank)):
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        losses.append(loss.item())
        saw_compressed |= not engine.enable_backward_allreduce
    assert saw_compressed, "the optimizer never left its warm-up stage"
    assert all(np.isfinite(losses))
    # every rank applies identical (all-reduced) updates: parameters must stay bit-identical across ranks
    if opt_type != "ZeroOneAdam":  # 0/1 Adam takes genuinely local steps between synchronisations
        flat = torch.cat([p.detach().float().reshape(-1) for p in engine.module.parameters()])
        ref = flat.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(flat, ref)
    assert losses[-1] < losses[0] * 1.5


@pytest.mark.parametrize("opt_type,params", [
    ("OnebitAdam", {"lr": 1e-2, "freeze_step": 3, "comm_backend_name": "nccl"}),
    ("OnebitLamb", {"lr": 1e-2, "freeze_step": 3, "comm_backend_name": "nccl", "coeff_beta": 0.9}),
    ("ZeroOneAdam", {"lr": 1e-2, "var_freeze_step": 4, "var_update_scaler": 2, "local_step_scaler": 2, "local_step_clipper": 2}),
])
def test_compressed_optimizers_train(opt_type, params):
    run_distributed(_train_body, world_size=2, args=(opt_type, params))
