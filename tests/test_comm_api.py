"""The comm layer's collectives against their definitions, 3 ranks over gloo (reference test strategy: tests/unit/comm/test_dist.py),
plus the comms logger collecting through the `timed_op` wrapper."""
import torch

from common import run_distributed


def _body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    r, W = dist.get_rank(), dist.get_world_size()
    assert dist.is_initialized() and dist.get_backend() == "gloo" and dist.get_all_ranks_from_group() == list(range(W))
    # all_reduce: sum / max / min / product / avg
    t = torch.full((5, ), float(r + 1))
    dist.all_reduce(t)
    assert torch.all(t == sum(range(1, W + 1)))
    for op, want in ((dist.ReduceOp.MAX, float(W)), (dist.ReduceOp.MIN, 1.0), (dist.ReduceOp.PRODUCT, 6.0)):
        t = torch.tensor([float(r + 1)])
        dist.all_reduce(t, op=op)
        assert t.item() == want
    # broadcast / reduce
    t = torch.arange(4.0) if r == 1 else torch.zeros(4)
    dist.broadcast(t, src=1)
    assert torch.equal(t, torch.arange(4.0))
    t = torch.tensor([float(r)])
    dist.reduce(t, dst=2)
    assert r != 2 or t.item() == sum(range(W))
    # all_gather in both forms, uneven coalesced gather
    out = torch.empty(W * 3)
    dist.all_gather_into_tensor(out, torch.full((3, ), float(r)))
    assert torch.equal(out, torch.arange(W).repeat_interleave(3).float())
    lst = [torch.empty(2) for _ in range(W)]
    dist.all_gather(lst, torch.full((2, ), float(r)))
    assert [x[0].item() for x in lst] == [0.0, 1.0, 2.0]
    # reduce_scatter in both forms
    out = torch.empty(2)
    dist.reduce_scatter_tensor(out, torch.arange(W * 2.0) + r)
    assert torch.equal(out, W * torch.arange(2 * r, 2 * r + 2.0) + sum(range(W)))
    out = torch.empty(1)
    dist.reduce_scatter(out, [torch.tensor([float(r * 10 + j)]) for j in range(W)])
    assert out.item() == sum(i * 10 + r for i in range(W))
    # all_to_all: equal and unequal splits
    out = torch.empty(W)
    dist.all_to_all_single(out, torch.arange(W).float() + 10 * r)
    assert torch.equal(out, torch.tensor([10.0 * i + r for i in range(W)]))
    send_counts = [r + 1] * W  # rank r sends r + 1 elements to everybody
    recv_counts = [i + 1 for i in range(W)]
    out = torch.empty(sum(recv_counts))
    dist.all_to_all_single(out, torch.full((sum(send_counts), ), float(r)), output_split_sizes=recv_counts, input_split_sizes=send_counts)
    assert torch.equal(out, torch.cat([torch.full((i + 1, ), float(i)) for i in range(W)]))
    outs = [torch.empty(1) for _ in range(W)]
    dist.all_to_all(outs, [torch.tensor([float(100 * r + j)]) for j in range(W)])
    assert [o.item() for o in outs] == [100.0 * i + r for i in range(W)]
    # point-to-point ring, blocking and non-blocking
    nxt, prv = (r + 1) % W, (r - 1) % W
    got = torch.empty(1)
    if r % 2 == 0:
        dist.send(torch.tensor([float(r)]), dst=nxt)
        dist.recv(got, src=prv)
    else:
        dist.recv(got, src=prv)
        dist.send(torch.tensor([float(r)]), dst=nxt)
    assert got.item() == prv
    got = torch.empty(1)
    reqs = [dist.isend(torch.tensor([float(r + 7)]), dst=nxt), dist.irecv(got, src=prv)]
    for q in reqs:
        q.wait()
    assert got.item() == prv + 7
    # gather / scatter / objects
    glist = [torch.empty(1) for _ in range(W)] if r == 0 else None
    dist.gather(torch.tensor([float(r)]), glist, dst=0)
    assert r != 0 or [g.item() for g in glist] == [0.0, 1.0, 2.0]
    t = torch.empty(1)
    dist.scatter(t, [torch.tensor([float(5 + j)]) for j in range(W)] if r == 0 else None, src=0)
    assert t.item() == 5 + r
    objs = [None] * W
    dist.all_gather_object(objs, {"rank": r})
    assert [o["rank"] for o in objs] == list(range(W))
    box = [("payload", 3)] if r == 0 else [None]
    dist.broadcast_object_list(box, src=0)
    assert box[0] == ("payload", 3)
    # sub-groups: collectives stay inside, ranks translate
    pair = dist.new_group([0, 2])
    solo = dist.new_group([1])
    if r in (0, 2):
        t = torch.tensor([float(r)])
        dist.all_reduce(t, group=pair)
        assert t.item() == 2.0 and dist.get_world_size(pair) == 2 and dist.get_global_rank(pair, 1) == 2 and dist.get_rank(pair) == r // 2
    else:
        assert dist.get_world_size(solo) == 1
    dist.barrier()
    # comms logger through the timed_op wrapper
    dist.configure(enabled=True, prof_all=True, verbose=False)
    t = torch.ones(256)
    for _ in range(3):
        dist.all_reduce(t, log_name="grad_sync")
    from deepspeedsyclsupport_b200.comm.comm import comms_logger
    rec = comms_logger.comms_dict["grad_sync"][256 * 4]
    assert rec[0] == 3 and len(rec[1]) == 3 and all(ms >= 0 for ms in rec[1])
    dist.log_summary()
    dist.configure(enabled=False)


def test_comm_collectives_three_ranks():
    run_distributed(_body, world_size=3)
