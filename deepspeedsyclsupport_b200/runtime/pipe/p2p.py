"""Pipeline-stage point-to-point transfers (async isend / irecv on NCCL, overlapped with compute).

Parity: deepspeed/runtime/pipe/p2p.py:50-163 -- same entry points (`init_process_groups`, `send`, `recv`, `wait`, `send_obj`,
`recv_obj`), addressed by pipeline STAGE; the reference asserts async off, here transfers may be asynchronous and the engine
waits right before the data is consumed.  Python objects travel as ONE framed byte tensor: an 8-byte little-endian length
followed by the pickle, padded to the next power-of-two bucket so the receiver can post its buffer without a size handshake
for small messages (one send instead of two)."""
import pickle
import struct

import torch

from ... import comm as dist
from ...accelerator import get_accelerator

_grid = None
_groups = None
_SMALL = 4096  # objects whose frame fits are sent in one fixed-size message


def can_send_recv() -> bool:
    """Point-to-point send/recv is always available on this stack (parity: runtime/pipe/p2p.py:23-26)."""
    return True


def init_process_groups(grid):
    global _grid
    assert grid.pipe_parallel_size > 1, "There is no pipeline parallelism"
    _grid = grid


def _peer_rank(other_stage):
    """Global rank of `other_stage` in this rank's pipeline; only neighbours (with wrap-around first <-> last) may talk."""
    here, stages = _grid.get_stage_id(), _grid.pipe_parallel_size
    assert (here - other_stage) % stages in (1, stages - 1), \
        "Functionality currently limited to send and receive between adjacent ranks only"
    return _grid.stage_to_global(stage_id=other_stage)


def send(tensor, dest_stage, async_op=False):
    rank = _peer_rank(dest_stage)
    return dist.isend(tensor, rank) if async_op else dist.send(tensor, rank)


def recv(tensor, src_stage, async_op=False):
    rank = _peer_rank(src_stage)
    return dist.irecv(tensor, src=rank) if async_op else dist.recv(tensor, src=rank)


def wait():
    get_accelerator().synchronize()


def _device():
    return get_accelerator().device_name()


def send_obj(msg, dest: int):
    """Pickle `msg` and send it to global rank `dest` (tensors inside travel by value)."""
    body = pickle.dumps(msg)
    frame = struct.pack("<Q", len(body)) + body
    head = torch.frombuffer(bytearray(frame[:_SMALL].ljust(_SMALL, b"\0")), dtype=torch.uint8).to(_device())
    dist.send(head, dst=dest)
    if len(frame) > _SMALL:  # the tail of a large object follows in a second message of exactly the missing size
        dist.send(torch.frombuffer(bytearray(frame[_SMALL:]), dtype=torch.uint8).to(_device()), dst=dest)


def _to_device(x):
    """Move every tensor inside a nested structure of tuples / lists / mappings to the accelerator."""
    if torch.is_tensor(x):
        return x.to(_device())
    if isinstance(x, (tuple, list)):
        return type(x)(_to_device(v) for v in x)
    if hasattr(x, "items"):
        out = x.__class__()
        for k, v in x.items():
            out[_to_device(k)] = _to_device(v)
        return out
    return x


def recv_obj(sender: int):
    head = torch.empty(_SMALL, dtype=torch.uint8, device=_device())
    dist.recv(head, src=sender)
    raw = head.cpu().numpy().tobytes()
    (size, ) = struct.unpack("<Q", raw[:8])
    body = raw[8:8 + size]
    if 8 + size > _SMALL:
        tail = torch.empty(8 + size - _SMALL, dtype=torch.uint8, device=_device())
        dist.recv(tail, src=sender)
        body += tail.cpu().numpy().tobytes()
    return _to_device(pickle.loads(body))
