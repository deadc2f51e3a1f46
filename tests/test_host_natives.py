"""Host natives: AVX-512 CPU optimizers, async file I/O, KV block allocator, atom builder."""
import ctypes
import os

import pytest
import torch


def test_cpu_adam_matches_torch():
    from deepspeedsyclsupport_b200.ops.adam import DeepSpeedCPUAdam
    torch.manual_seed(0)
    w = [torch.randn(1000, 17, requires_grad=True), torch.randn(333, requires_grad=True)]
    r = [t.detach().clone().requires_grad_() for t in w]
    o, ro = DeepSpeedCPUAdam(w, lr=1e-2, weight_decay=0.05), torch.optim.AdamW(r, lr=1e-2, weight_decay=0.05)
    for _ in range(4):
        for a, b in zip(w, r):
            g = torch.randn_like(a)
            a.grad, b.grad = g.clone(), g.clone()
        o.step()
        ro.step()
    for a, b in zip(w, r):
        assert torch.allclose(a, b, atol=2e-6)


def test_cpu_adam_writes_bf16_copy():
    from deepspeedsyclsupport_b200.ops.flat_optim import adam_step
    p, m, v = torch.randn(5000), torch.zeros(5000), torch.zeros(5000)
    g = torch.randn(5000).bfloat16()
    lp = torch.empty(5000, dtype=torch.bfloat16)
    adam_step(p, m, v, g, lr=1e-2, step=1, lp=lp)
    assert torch.equal(lp, p.bfloat16())


def test_cpu_lion_and_adagrad():
    from deepspeedsyclsupport_b200.ops.adagrad import DeepSpeedCPUAdagrad
    from deepspeedsyclsupport_b200.ops.lion import DeepSpeedCPULion
    torch.manual_seed(0)
    w = torch.randn(4096, requires_grad=True)
    w0, g = w.detach().clone(), torch.randn(4096)
    w.grad = g.clone()
    DeepSpeedCPULion([w], lr=1e-2, betas=(0.9, 0.99)).step()
    assert torch.allclose(w.detach(), w0 - 1e-2 * torch.sign(0.1 * g), atol=1e-6)
    w = torch.randn(4096, requires_grad=True)
    r = w.detach().clone().requires_grad_()
    o, ro = DeepSpeedCPUAdagrad([w], lr=1e-2, eps=1e-10), torch.optim.Adagrad([r], lr=1e-2, eps=1e-10)
    for _ in range(3):
        g = torch.randn(4096)
        w.grad, r.grad = g.clone(), g.clone()
        o.step()
        ro.step()
    assert torch.allclose(w, r, atol=1e-5)


def test_aio_roundtrip(tmp_path):
    from deepspeedsyclsupport_b200.ops.aio import AsyncIOBuilder
    h = AsyncIOBuilder().load().aio_handle(block_size=64 * 1024, queue_depth=4, single_submit=False, overlap_events=True, num_threads=2)
    a = torch.randn(1 << 18)
    f = str(tmp_path / "t.swp")
    h.sync_pwrite(a, f)
    b = torch.empty_like(a)
    h.sync_pread(b, f)
    assert torch.equal(a, b)
    files = [str(tmp_path / f"x{i}.swp") for i in range(4)]
    srcs = [torch.randn(100_003) for _ in files]
    for s, p in zip(srcs, files):
        h.async_pwrite(s, p)
    assert h.wait() == 4
    outs = [torch.empty(100_003) for _ in files]
    for o, p in zip(outs, files):
        h.async_pread(o, p)
    h.wait()
    assert all(torch.equal(s, o) for s, o in zip(srcs, outs))
    with pytest.raises(IOError):
        h.sync_pread(torch.empty(10), str(tmp_path / "missing.swp"))


def test_optimizer_state_swapper(tmp_path):
    from deepspeedsyclsupport_b200.runtime.swap_tensor import OptimizerStateSwapper
    sw = OptimizerStateSwapper(tmp_path, {"block_size": 1 << 16, "queue_depth": 4, "thread_count": 1}, [1000, 2000])
    st = [{n: torch.randn(k) for n in ("master", "exp_avg", "exp_avg_sq")} for k in (1000, 2000)]
    for g, s in enumerate(st):
        sw.initialize(g, s)
    sw.prefetch(0)
    a = sw.swap_in(0)
    sw.prefetch(1)
    assert torch.equal(a["master"], st[0]["master"])
    a["master"].add_(1.0)
    sw.swap_out(0)
    b = sw.swap_in(1)
    assert torch.equal(b["exp_avg_sq"], st[1]["exp_avg_sq"])
    sw.flush()
    assert torch.equal(sw.swap_in(0)["master"], st[0]["master"] + 1.0)


def test_blocked_allocator_and_atoms():
    from deepspeedsyclsupport_b200.inference.v2.ragged.blocked_allocator import BlockedAllocator
    a = BlockedAllocator(16)
    x = a.allocate(5)
    assert a.free_blocks == 11 and len(set(x.tolist())) == 5
    with pytest.raises(ValueError):
        a.allocate(12)
    a.free(x[:2])
    assert a.free_blocks == 13
    with pytest.raises(ValueError):
        a.free(x[:1])  # double free
    from deepspeedsyclsupport_b200.inference.v2.ragged.ragged_wrapper import build_atoms
    atoms = build_atoms(seq_start=[0, 300], seq_ntok=[300, 1], seq_seen=[0, 77], q_block=128, kv_block=64)
    assert atoms.shape == (4, 6)
    assert atoms[2].tolist() == [0, 256, 44, 5, 300, 256]
    assert atoms[3].tolist() == [1, 300, 1, 2, 78, 77]


def _shm_body():
    import torch
    import deepspeedsyclsupport_b200.comm as dist
    from deepspeedsyclsupport_b200.comm import shm
    rank, world = dist.get_rank(), dist.get_world_size()
    seg = shm.for_group(None)
    assert seg is not None and seg.world == world
    for dtype, n in ((torch.float32, 1000), (torch.bfloat16, 3 * (1 << 20) + 17), (torch.float16, 33), (torch.int32, 5)):
        g = torch.Generator().manual_seed(5)
        parts = [(torch.randn(n, generator=g) * 4).to(dtype) if dtype != torch.int32 else torch.arange(n, dtype=dtype) * (r + 1)
                 for r in range(world)]
        want = parts[0].float()
        for p in parts[1:]:
            want = want + p.float()
        t = parts[rank].clone()
        assert dist.inference_all_reduce(t) is None
        tol = {torch.float32: 1e-6, torch.bfloat16: 2e-2, torch.float16: 2e-3, torch.int32: 0}[dtype]
        assert (t.float() - want).abs().max() <= tol * max(1.0, want.abs().max().item())
    # a non-contiguous tensor falls back to the backend collective and still sums
    x = torch.ones(4, 6)[:, ::2] * (rank + 1)
    dist.inference_all_reduce(x)
    assert torch.all(x == sum(range(1, world + 1)))
    seg.barrier()


def test_shm_all_reduce_three_ranks():
    """POSIX-shm all-reduce (csrc/host/shm_comm.cpp): several dtypes, multi-chunk sizes, fallback for strided tensors."""
    from common import run_distributed
    run_distributed(_shm_body, world_size=3)


def test_sparse_lut_matches_nonzero():
    import torch
    from deepspeedsyclsupport_b200.ops.sparse_attention.matmul import build_lut
    torch.manual_seed(0)
    layout = (torch.rand(3, 7, 9) < 0.3).long()
    h, r, c, rp = build_lut(layout)
    nz = layout.nonzero()
    assert torch.equal(torch.stack([h, r, c], 1), nz)
    counts = layout.view(21, 9).sum(1)
    assert torch.equal(rp[1:] - rp[:-1], counts) and rp[-1] == nz.shape[0]


def test_partitioned_param_swapper_roundtrip(tmp_path):
    """NVMe tier for parameter shards: swap out (async), read back ahead of use, values intact, buffers recycled."""
    import torch
    from deepspeedsyclsupport_b200.runtime.swap_tensor import AsyncPartitionedParameterSwapper, PartitionedParamStatus
    cfg = {"zero_optimization": {"offload_param": {"device": "nvme", "nvme_path": str(tmp_path), "buffer_count": 3, "buffer_size": 5000}},
           "aio": {"block_size": 65536, "queue_depth": 4}}
    sw = AsyncPartitionedParameterSwapper(cfg, torch.bfloat16)
    params, want = [], []
    for i in range(5):
        p = torch.nn.Parameter(torch.empty(0))
        p.ds_id = i
        p.ds_tensor = torch.randn(1000 + 37 * i).bfloat16()
        want.append(p.ds_tensor.clone())
        params.append(p)
    for p in params:
        assert sw.swappable_tensor(p)
        sw.swap_out_and_release([p])
        assert p.ds_tensor.numel() == 1 and p.ds_tensor.status == PartitionedParamStatus.NOT_AVAILABLE
    assert sw.available_swap_in_buffers() == 3
    sw.swap_in(params[:3], async_op=True)
    assert sw.available_swap_in_buffers() == 0 and params[0].ds_tensor.status == PartitionedParamStatus.INFLIGHT
    sw.synchronize_reads()
    for p, w in zip(params[:3], want[:3]):
        assert p.ds_tensor.status == PartitionedParamStatus.AVAILABLE and torch.equal(p.ds_tensor, w)
    sw.remove_partition_and_release_buffers(params[:3])
    sw.swap_in(params[3:], async_op=False)
    assert torch.equal(params[4].ds_tensor, want[4])


def _ccl_body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.comm.ccl import CCLBackend, CCLHandler
    be = CCLBackend(rank=dist.get_rank(), world_size=dist.get_world_size())
    assert be.is_initialized() and be.name == "ccl" and "inference_all_reduce" in be.get_all_available_coll()
    r, W = dist.get_rank(), dist.get_world_size()
    t = torch.full((1000, ), float(r + 1), dtype=torch.bfloat16)
    h = be.inference_all_reduce(t)
    assert isinstance(h, CCLHandler) and h.is_completed()  # went through the shared-memory segment
    assert torch.all(t == sum(range(1, W + 1)))
    m = torch.tensor([float(r)])
    be.all_reduce(m, op=dist.ReduceOp.MAX)  # not a sum: gloo
    assert m.item() == W - 1
    be.run_collective("all_reduce", tensor=t)
    assert torch.all(t == W * sum(range(1, W + 1)))


def test_ccl_backend_uses_shared_memory_for_sums():
    from common import run_distributed
    run_distributed(_ccl_body, world_size=2)


def test_comm_utils_find_the_payload(monkeypatch):
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.comm import utils as U
    from deepspeedsyclsupport_b200.comm.constants import COMMS_LOGGER_PROF_ALL_DEFAULT, NCCL_BACKEND
    from deepspeedsyclsupport_b200.constants import TORCH_DISTRIBUTED_DEFAULT_PORT, default_pg_timeout

    def all_reduce(tensor, op=None, group=None, async_op=False):
        pass

    def all_gather(tensor_list, tensor, group=None):
        pass

    def coalesced(tensors, group=None):
        pass

    t = torch.zeros(10, dtype=torch.bfloat16)
    assert U.get_tensor_position(all_reduce) == 0 and U.get_tensor_position(all_gather) == 1
    assert U.get_msg_size_from_args(all_reduce, t) == 20 and U.get_msg_size_from_args(all_reduce, tensor=t) == 20
    assert U.get_msg_size_from_args(coalesced, [t, t.float()]) == 60
    assert U.get_default_args(all_reduce) == {"op": None, "group": None, "async_op": False}
    assert U.get_debug_log_name({"log_name": "x"}, "caller") == "x | [Caller Func: caller]" and U.get_debug_log_name({"log_name": "x"}, None) == "x"
    for k in ("LOCAL_RANK", "RANK", "WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID", "OMPI_COMM_WORLD_RANK", "SLURM_PROCID",
              "OMPI_COMM_WORLD_SIZE", "SLURM_NPROCS"):
        monkeypatch.delenv(k, raising=False)
    assert (U.get_local_rank_from_launcher(), U.get_world_rank_from_launcher(), U.get_world_size_from_launcher()) == (0, 0, 1)
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "5")
    monkeypatch.setenv("SLURM_NPROCS", "8")
    assert U.get_world_rank_from_launcher() == 5 and U.get_world_size_from_launcher() == 8
    assert NCCL_BACKEND == "nccl" and COMMS_LOGGER_PROF_ALL_DEFAULT is True and TORCH_DISTRIBUTED_DEFAULT_PORT == 29500
    assert default_pg_timeout.total_seconds() > 0 and dist.ReduceOp.SUM is not None


def test_async_tensor_swapper_and_gradient_bookkeeping(tmp_path):
    from deepspeedsyclsupport_b200.ops.aio import aio_handle
    from deepspeedsyclsupport_b200.runtime.swap_tensor.async_swapper import AsyncTensorSwapper
    from deepspeedsyclsupport_b200.runtime.swap_tensor.optimizer_utils import OptimizerStateSwapInfo, OptimizerSwapper
    h = aio_handle(block_size=4096, queue_depth=4)
    sw = AsyncTensorSwapper(h, numel_alignment=256)
    sw.add_buffers([torch.zeros(1024), torch.zeros(1024)])
    torch.manual_seed(0)
    tensors = [torch.randn(n) for n in (300, 500, 256, 900, 100, 700)]  # several buffers' worth: forces waits and reuse
    paths = [str(tmp_path / f"t{i}.swp") for i in range(len(tensors))]
    sw.swap_out_tensors(tensors, paths)
    buffers = sw.release_buffers()
    assert len(buffers) == 2 and not sw.has_buffers()
    for t, p in zip(tensors, paths):
        back = torch.empty(sw._io_aligned_numel(t.numel()))
        h.sync_pread(back, p)
        assert torch.equal(back[:t.numel()], t)
    # optimizer-side bookkeeping: big aligned gradient slices go to disk, small ones stay in memory, both come back in place
    p = torch.nn.Parameter(torch.zeros(4096))
    opt = torch.optim.Adam([p])
    osw = OptimizerSwapper(None, {"block_size": 1024, "thread_count": 1}, str(tmp_path), opt, largest_numel=4096, device="cpu", dtype=torch.float32)
    osw.min_aio_bytes = 1024  # tiny threshold for the test
    info = osw._create_param_swap_info(p, p.numel())
    assert isinstance(info, OptimizerStateSwapInfo) and osw.swappable_tensor(numel=256) and not osw.swappable_tensor(numel=100)
    grads = {0: torch.randn(1024), 1024: torch.randn(100), 2048: torch.randn(768)}
    gsw = AsyncTensorSwapper(h, osw.numel_alignment)
    gsw.add_buffers(buffers)
    osw._swap_out_gradients(p, list(grads), list(grads.values()), gsw)
    assert sorted(info.swapped_gradients) == [0, 2048] and sorted(info.unswapped_gradients) == [1024]
    osw._flush_gradient_swapper(gsw)
    dest = torch.zeros(4096)
    n = osw._swap_in_gradients(h, p, dest)
    assert n == 1024 + 100 + 768
    for off, g in grads.items():
        assert torch.equal(dest[off:off + g.numel()], g)
    assert not info.unswapped_gradients
