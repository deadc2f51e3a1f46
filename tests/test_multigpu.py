"""Multi-GPU tests (NCCL + peer memory).  Run on a box with >= 2 GPUs: pytest -m gpu tests/test_multigpu.py"""
import pytest
import torch

from common import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _symm_body():
    import deepspeedsyclsupport_b200 as ds
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.parallel.symm import SymmetricBuffer
    rank, W = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    n = 8 * 4096
    sb = SymmetricBuffer(dist.get_world_group(), W * n * 4)
    torch.manual_seed(100 + rank)
    grad = torch.randn(W * n, device=dev).bfloat16()
    sb.view(0, (W * n, ), torch.bfloat16).copy_(grad)
    sb.barrier()
    dst = torch.zeros(n, device=dev)
    sb.reduce_scatter(dst, 0, torch.bfloat16, n, scale=1.0 / W)
    ref = grad.float().clone()
    dist.all_reduce(ref)
    ref = ref[rank * n:(rank + 1) * n] / W
    torch.cuda.synchronize()
    assert (dst - ref).abs().max() < 1e-5, (dst - ref).abs().max()
    sb.barrier()
    # all-gather of the first n elements of every rank
    out = torch.empty(W * n, device=dev, dtype=torch.bfloat16)
    sb.all_gather(out, 0, n * 2)
    ref2 = torch.empty(W * n, device=dev, dtype=torch.bfloat16)
    dist.all_gather_into_tensor(ref2, grad[:n].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(out, ref2)
    sb.barrier()
    # all-reduce, one-shot and two-shot (reduce-scatter + barrier + all-gather through a symmetric scratch slice)
    m = 16 * W * 1024
    sb2 = SymmetricBuffer(dist.get_world_group(), m * 2 + m * 2 // W + 256)
    v = torch.randn(m, device=dev).bfloat16()
    sb2.view(0, (m, ), torch.bfloat16).copy_(v)
    sb2.barrier()
    want = v.float().clone()
    dist.all_reduce(want)
    for algo in ("one_shot", "two_shot"):
        got = torch.empty(m, device=dev, dtype=torch.bfloat16)
        sb2.all_reduce_bf16(got, 0, m, scratch_offset=m * 2, algo=algo)
        sb2.barrier()
        torch.cuda.synchronize()
        assert (got.float() - want).abs().max() <= 2e-2 * want.abs().max(), algo
    sb2.close()
    torch.cuda.synchronize()
    sb.close()


def test_symmetric_memory_collectives_match_nccl():
    run_distributed(_symm_body, world_size=2, backend="nccl")


def _zero3_body():
    import copy
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    rank = dist.get_rank()
    torch.manual_seed(0)
    cfg = llama_config("tiny", hidden_size=256, intermediate_size=768, num_attention_heads=2, num_key_value_heads=2,
                       num_hidden_layers=3, vocab_size=1024, activation_checkpointing=True)
    model = LlamaForCausalLM(cfg).cuda().bfloat16()
    engine, _, _, _ = deepspeed.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "gradient_clipping": 1.0,
        "optimizer": {"type": "AdamW", "params": {"lr": 2e-3}},
        "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}})
    g = torch.Generator().manual_seed(5 + rank)
    ids = torch.randint(0, 1024, (2, 128), generator=g).cuda()
    labels = torch.randint(0, 1024, (2, 128), generator=g).cuda()
    losses = []
    for _ in range(8):
        loss = engine(ids, labels)
        engine.backward(loss)
        engine.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0] - 0.5, losses


def test_zero3_two_gpus_trains():
    run_distributed(_zero3_body, world_size=2, backend="nccl")
