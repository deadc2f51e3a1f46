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


def _peer_a2a_body():
    """comm.all_to_all_single on device tensors takes the peer-memory pull kernel inside one node and must equal NCCL; the
    Ulysses exchange built on it must round-trip; timing of both paths is printed (device events, max over ranks)."""
    import os
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.parallel.symm import PeerAllToAll
    from deepspeedsyclsupport_b200.sequence.layer import single_all_to_all
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(rank)
    for n in (world * 1024, world * (1 << 20)):
        x = torch.randn(n, device="cuda").bfloat16()
        want = torch.empty_like(x)
        os.environ["DSB200_PEER_A2A"] = "0"
        dist.all_to_all_single(want, x)
        os.environ["DSB200_PEER_A2A"] = "1"
        got = torch.empty_like(x)
        dist.all_to_all_single(got, x)
        assert torch.equal(got, want)
    assert PeerAllToAll._instances, "the peer path was not taken"
    q = torch.randn(64, 2, 4 * world, 32, device="cuda").bfloat16()  # [s/P, b, h, d]
    full = single_all_to_all(q, 2, 0, None)
    assert full.shape == (64 * world, 2, 4, 32)
    back = single_all_to_all(full, 0, 2, None)
    assert torch.equal(back, q)

    def timed(flag):
        os.environ["DSB200_PEER_A2A"] = flag
        x = torch.randn(world * (4 << 20), device="cuda").bfloat16()
        y = torch.empty_like(x)
        for _ in range(3):
            dist.all_to_all_single(y, x)
        torch.cuda.synchronize()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(10):
            dist.all_to_all_single(y, x)
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e) / 10], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    t_nccl, t_peer = timed("0"), timed("1")
    if rank == 0:
        mb = world * (4 << 20) * 2 / 2**20
        print(f"ALLTOALL {mb:.0f} MiB per rank, world {world}: nccl {t_nccl:.3f} ms, peer {t_peer:.3f} ms, speedup {t_nccl / t_peer:.2f}x", flush=True)
    os.environ["DSB200_PEER_A2A"] = "1"


def test_peer_memory_all_to_all_matches_nccl():
    run_distributed(_peer_a2a_body, world_size=2, backend="nccl")


def _gemm_ar_body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    from deepspeedsyclsupport_b200.parallel.fused import FusedGemmAllReduce
    from deepspeedsyclsupport_b200.parallel.symm import SymmetricBuffer
    rank, W = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    for (M, N, K) in [(200, 1024, 512), (2048, 8192, 4096)]:
        sb = SymmetricBuffer(dist.get_world_group(), FusedGemmAllReduce.bytes_needed(M, N))
        op = FusedGemmAllReduce(sb, M, N)
        sb.barrier()
        torch.cuda.synchronize()
        dist.barrier()
        torch.manual_seed(7 + rank)
        x = torch.randn(M, K, device=dev).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.03).bfloat16()
        torch.manual_seed(3)
        res = torch.randn(M, N, device=dev).bfloat16()
        want = x.float() @ w.float().t()
        dist.all_reduce(want)
        for it in range(4):  # both partial areas, increasing epochs
            r = res if it % 2 else None
            got = op(x, w, residual=r).float()
            torch.cuda.synchronize()
            ref = want + (r.float() if r is not None else 0)
            err = (got - ref).abs().max().item()
            assert err <= 2e-2 * ref.abs().max().item(), (M, N, K, it, err)

        def t(fn, n=10):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(n):
                fn()
            e.record()
            torch.cuda.synchronize()
            v = torch.tensor([s.elapsed_time(e) / n], device=dev)
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            return v.item()

        def base():
            y = gemm(x, w)
            dist.all_reduce(y)
            return y

        if M % W == 0:  # reduce-scatter form: each rank keeps its slice of the rows
            from deepspeedsyclsupport_b200.parallel.fused import FusedGemmReduceScatter
            rs = FusedGemmReduceScatter.__new__(FusedGemmReduceScatter)
            rs.__dict__.update(op.__dict__)  # same symmetric areas and epoch counter as `op`
            lo, hi = rank * M // W, (rank + 1) * M // W
            for it in range(3):
                got = rs(x, w, residual=res[lo:hi] if it == 1 else None).float()
                op.epoch = rs.epoch
                torch.cuda.synchronize()
                ref = want[lo:hi] + (res[lo:hi].float() if it == 1 else 0)
                err = (got - ref).abs().max().item()
                assert got.shape == (hi - lo, N) and err <= 2e-2 * ref.abs().max().item(), ("rs", M, N, K, it, err)

            def base_rs():
                y = gemm(x, w)
                o = torch.empty(M // W, N, dtype=y.dtype, device=dev)
                dist.reduce_scatter_tensor(o, y)
                return o

            t_rs, t_rs_base = t(lambda: rs(x, w)), t(base_rs)
            op.epoch = rs.epoch
            if rank == 0:
                print(f"GEMM_RS M={M} N={N} K_local={K} world {W}: gemm+nccl {t_rs_base:.3f} ms, fused {t_rs:.3f} ms, "
                      f"speedup {t_rs_base / t_rs:.2f}x", flush=True)
        t_fused, t_base = t(lambda: op(x, w)), t(base)
        if rank == 0:
            print(f"GEMM_AR M={M} N={N} K_local={K} world {W}: gemm+nccl {t_base:.3f} ms, fused {t_fused:.3f} ms, "
                  f"speedup {t_base / t_fused:.2f}x", flush=True)
        sb.barrier()
        torch.cuda.synchronize()
        sb.close()


def test_fused_gemm_allreduce_matches_gemm_plus_nccl():
    run_distributed(_gemm_ar_body, world_size=2, backend="nccl")


def _tp_fused_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.module_inject.layers import LinearAllreduce
    from test_inference_v1 import TinyDecoderModel
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(0)
    model = TinyDecoderModel(h=256, heads=4, n=2, vocab=64).eval()
    ids = torch.randint(0, 64, (2, 40))
    with torch.no_grad():
        ref = model(ids).float()
    engine = deepspeed.init_inference(model, tensor_parallel={"tp_size": 2}, dtype=torch.bfloat16)
    rows = [m for m in engine.module.modules() if isinstance(m, LinearAllreduce)]
    assert len(rows) == 4 and all(m._fused is not None for m in rows), "fused GEMM + all-reduce not attached"
    with torch.no_grad():
        for _ in range(3):  # repeated forwards: epochs / both partial areas
            out = engine(ids.to(dev)).float().cpu()
    assert rows[0]._fused.epoch == 12
    assert (out - ref).abs().max() <= 0.06 * ref.abs().max(), (out - ref).abs().max()


def test_init_inference_tp_uses_fused_gemm_allreduce():
    run_distributed(_tp_fused_body, world_size=2, backend="nccl")
