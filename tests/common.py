"""Single-host multi-process test harness (the idea of the reference's tests/unit/common.py:105-237 DistributedExec):
spawn `world_size` ranks on this machine, rendezvous on 127.0.0.1 with a free port, run the test body in every rank,
propagate failures/skips, and treat a hang as a failure (timeout)."""
import multiprocessing as mp
import os
import socket
import sys
import traceback

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, backend, fn, args, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world_size), LOCAL_SIZE=str(world_size), LOCAL_WORLD_SIZE=str(world_size))
    if backend == "gloo":
        os.environ["DS_ACCELERATOR"] = "cpu"
    sys.path.insert(0, ROOT)
    try:
        import torch
        torch.set_num_threads(1)
        import deepspeedsyclsupport_b200 as ds
        ds.init_distributed(dist_backend=backend)
        fn(*args)
        ds.comm.barrier()
        q.put((rank, "ok", ""))
    except pytest.skip.Exception as e:
        q.put((rank, "skip", str(e)))
    except BaseException:
        q.put((rank, "fail", traceback.format_exc()))
    finally:
        try:
            import torch.distributed as d
            if d.is_initialized():
                d.destroy_process_group()
        except Exception:
            pass


def run_distributed(fn, world_size=2, backend="gloo", args=(), timeout=240):
    """Run fn(*args) in `world_size` fresh processes (spawn).  `fn` must be a module-level function."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, backend, fn, args, q)) for r in range(world_size)]
    for p in procs:
        p.start()
    results = []
    try:
        for _ in range(world_size):
            results.append(q.get(timeout=timeout))
    except Exception:
        for p in procs:
            if p.is_alive():
                p.terminate()
        pytest.fail(f"distributed test hung or crashed (got {len(results)}/{world_size} results): {results}")
    for p in procs:
        p.join(timeout=30)
        if p.is_alive():
            p.terminate()
    fails = [r for r in results if r[1] == "fail"]
    if fails:
        pytest.fail("\n".join(f"[rank {r}] {tb}" for r, _, tb in fails))
    if any(r[1] == "skip" for r in results):
        pytest.skip(results[0][2])
