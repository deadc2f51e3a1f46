"""Pipeline-stage point-to-point transfers on NCCL (async isend/irecv, overlapped with compute).
Parity: deepspeed/runtime/pipe/p2p.py:50-163 (the reference asserts async off; here transfers are async and the
engine waits right before the data is consumed)."""
import pickle

import torch

from ... import comm as dist
from ...accelerator import get_accelerator

_groups = None
_grid = None


def can_send_recv() -> bool:
    """Point-to-point send/recv is always available on this stack (parity: runtime/pipe/p2p.py:23-26)."""
    return True


def init_process_groups(grid):
    global _groups, _grid
    _grid = grid
    assert _grid.pipe_parallel_size > 1, "There is no pipeline parallelism"


def _is_valid_send_recv(src_stage, dest_stage):
    first_stage = 0
    last_stage = _grid.pipe_parallel_size - 1
    assert abs(src_stage - dest_stage) == 1 or (src_stage == first_stage and dest_stage == last_stage) or \
        (src_stage == last_stage and dest_stage == first_stage), "Functionality currently limited to send and receive between adjacent ranks only"


def send(tensor, dest_stage, async_op=False):
    src_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    dest_rank = _grid.stage_to_global(stage_id=dest_stage)
    if async_op:
        return dist.isend(tensor, dest_rank)
    return dist.send(tensor, dest_rank)


def recv(tensor, src_stage, async_op=False):
    dest_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    src_rank = _grid.stage_to_global(stage_id=src_stage)
    if async_op:
        return dist.irecv(tensor, src=src_rank)
    return dist.recv(tensor, src=src_rank)


def wait():
    get_accelerator().synchronize()


def send_obj(msg, dest: int):
    msg = pickle.dumps(msg)
    msg = torch.ByteTensor(torch.ByteStorage.from_buffer(msg)).to(get_accelerator().device_name())
    length_tensor = torch.tensor([len(msg)], dtype=torch.long).to(get_accelerator().device_name())
    dist.send(length_tensor, dst=dest)
    dist.send(msg, dst=dest)


def recv_obj(sender: int):
    length = torch.tensor([0], dtype=torch.long).to(get_accelerator().device_name())
    dist.recv(length, src=sender)
    msg = torch.empty(length.item(), dtype=torch.uint8).to(get_accelerator().device_name())
    dist.recv(msg, src=sender)
    msg = pickle.loads(msg.cpu().numpy().tobytes())

    def _to(x):
        if torch.is_tensor(x):
            return x.to(get_accelerator().device_name())
        if isinstance(x, (tuple, list)):
            ret = [_to(x_) for x_ in x]
            if isinstance(x, tuple):
                ret = tuple(ret)
            return ret
        if hasattr(x, "items"):
            ret = x.__class__()
            for key, val in x.items():
                ret[_to(key)] = _to(val)
            return ret
        return x

    return _to(msg)
