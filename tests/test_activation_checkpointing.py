"""deepspeed.checkpointing on the host (reference: tests/unit/runtime/activation_checkpointing/test_activation_checkpointing.py):
gradients equal the un-checkpointed run, dropout replays with the saved RNG state, non-tensor and multiple outputs pass through,
cpu_checkpointing keeps no activation on the compute device, partition_activations splits saved inputs over the model-parallel group."""
import pytest
import torch

from common import run_distributed


def _fresh(**kw):
    from deepspeedsyclsupport_b200.runtime.activation_checkpointing import checkpointing as C
    C.reset()
    C.configure(None, **kw)
    return C


class Block(torch.nn.Module):

    def __init__(self, p=0.0):
        super().__init__()
        self.a, self.b, self.drop = torch.nn.Linear(16, 32), torch.nn.Linear(32, 16), torch.nn.Dropout(p)

    def forward(self, x, scale=1.0):
        return x + self.b(self.drop(torch.relu(self.a(x)))) * scale


def _grads(model, fn, x):
    model.zero_grad()
    x = x.clone().requires_grad_(True)
    out = fn(x)
    out.pow(2).sum().backward()
    return out.detach(), x.grad.clone(), [p.grad.clone() for p in model.parameters()]


@pytest.mark.parametrize("cpu_checkpointing", [False, True])
def test_checkpoint_matches_plain_autograd(cpu_checkpointing):
    C = _fresh(partition_activations=False, checkpoint_in_cpu=cpu_checkpointing)
    torch.manual_seed(0)
    model, x = Block(), torch.randn(6, 16)
    ref = _grads(model, lambda t: model(model(t)), x)
    got = _grads(model, lambda t: C.checkpoint(model, C.checkpoint(model, t)), x)
    assert torch.allclose(ref[0], got[0]) and torch.allclose(ref[1], got[1], atol=1e-6)
    assert all(torch.allclose(a, b, atol=1e-6) for a, b in zip(ref[2], got[2]))


def test_checkpoint_replays_dropout_and_passes_non_tensors():
    C = _fresh(partition_activations=False)
    torch.manual_seed(0)
    model, x = Block(p=0.5), torch.randn(8, 16)
    model.train()
    torch.manual_seed(123)
    ref = _grads(model, lambda t: model(t, 2.0), x)
    torch.manual_seed(123)
    got = _grads(model, lambda t: C.checkpoint(model, t, 2.0), x)  # a float argument goes through untouched
    assert torch.allclose(ref[0], got[0]) and torch.allclose(ref[1], got[1], atol=1e-6)  # same dropout mask in forward and recompute
    assert all(torch.allclose(a, b, atol=1e-6) for a, b in zip(ref[2], got[2]))

    def two_outputs(t):
        h = model.a(t)
        return h, h.sum(-1), 7  # tensors and a non-tensor

    model.eval()
    h, s, k = C.checkpoint(two_outputs, x.clone().requires_grad_(True))
    assert h.shape == (8, 32) and s.shape == (8, ) and k == 7 and h.requires_grad
    (h.sum() + s.sum()).backward()
    assert model.a.weight.grad is not None


def test_rng_tracker_forks_are_reproducible_and_isolated():
    from deepspeedsyclsupport_b200.runtime.activation_checkpointing.checkpointing import get_cuda_rng_tracker
    tr = get_cuda_rng_tracker()
    tr.reset()
    tr.add("model-parallel-rng", 1234)
    with pytest.raises(Exception):
        tr.add("model-parallel-rng", 99)  # name taken
    with pytest.raises(Exception):
        tr.add("other", 1234)  # seed taken
    torch.manual_seed(5)
    before = torch.rand(3)
    torch.manual_seed(5)
    with tr.fork():
        inside1 = torch.rand(3)
    after = torch.rand(3)
    assert torch.equal(before, after)  # the fork did not advance the default stream
    saved = tr.get_states()
    with tr.fork():
        inside2 = torch.rand(3)
    assert not torch.equal(inside1, inside2)  # the tracked stream advances across forks
    tr.set_states(saved)
    with tr.fork():
        assert torch.equal(torch.rand(3), inside2)
    tr.reset()


class _MPU:
    """The two ranks form one model-parallel group."""

    def __init__(self, group):
        self.group = group

    def get_model_parallel_group(self):
        return self.group

    def get_model_parallel_rank(self):
        import torch.distributed as d
        return d.get_rank(self.group)

    def get_model_parallel_world_size(self):
        import torch.distributed as d
        return d.get_world_size(self.group)

    def get_data_parallel_group(self):
        return None

    def get_data_parallel_rank(self):
        return 0

    def get_data_parallel_world_size(self):
        return 1


def _partition_body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.runtime.activation_checkpointing import checkpointing as C
    C.reset()
    C.configure(_MPU(dist.get_world_group()), partition_activations=True)
    torch.manual_seed(0)  # both ranks hold the same activations (tensor parallel ranks see the full sequence)
    model, x = Block(), torch.randn(6, 16)
    ref = _grads(model, lambda t: model(model(t)), x)
    parts = C.partition_activations((x, ))
    assert parts[0][0].numel() == x.numel() // 2  # each rank keeps half of the saved input
    got = _grads(model, lambda t: C.checkpoint(model, C.checkpoint(model, t)), x)
    assert torch.allclose(ref[0], got[0]) and torch.allclose(ref[1], got[1], atol=1e-6)
    assert all(torch.allclose(a, b, atol=1e-6) for a, b in zip(ref[2], got[2]))
    C.reset()


def test_partitioned_activations_two_ranks():
    run_distributed(_partition_body, world_size=2)
