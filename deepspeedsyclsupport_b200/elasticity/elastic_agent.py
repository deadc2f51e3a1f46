"""DSElasticAgent: per-node supervisor for elastic training.

Responsibilities (the role of torchelastic's LocalElasticAgent subclass in the reference,
deepspeed/elasticity/elastic_agent.py:28-188, re-built here without private torch.distributed.elastic APIs):
  * start one worker process per local GPU with RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* /
    TORCHELASTIC_RESTART_COUNT / TORCHELASTIC_MAX_RESTARTS set, on top of the DeepSpeed launcher environment;
  * monitor them; when a worker fails, stop the local group and -- while restarts remain -- start a new attempt;
  * membership: every agent heartbeats into a shared TCPStore.  When the set of live nodes changes (a node died, or a
    new one is waiting and the count stays within [min_nodes, max_nodes]) the lowest-ranked live agent publishes a new
    membership epoch; every agent seeing a new epoch restarts its workers with ranks / world size recomputed from the
    member list.  Workers resume from the latest checkpoint; the elastic batch config (elasticity.compute_elastic_config)
    guarantees the global batch size stays valid for the new GPU count.
The store is hosted by the agent of node 0 on master_port + 1."""
import copy
import os
import signal
import subprocess
import sys
import time
from datetime import timedelta
from typing import Dict, List, Optional

from ..utils.logging import logger

HEARTBEAT_INTERVAL = 2.0
HEARTBEAT_TIMEOUT = 15.0


def assign_ranks(members: List[int], node_id: int, nproc_per_node: int):
    """(global ranks of this node's workers, world size) for a sorted member list."""
    members = sorted(members)
    pos = members.index(node_id)
    return [pos * nproc_per_node + i for i in range(nproc_per_node)], len(members) * nproc_per_node


def pick_members(live: List[int], min_nodes: int, max_nodes: int) -> Optional[List[int]]:
    """Members of the next epoch: the lowest max_nodes live node ids, or None while fewer than min_nodes are alive."""
    live = sorted(live)
    if len(live) < max(1, min_nodes):
        return None
    return live[:max_nodes] if max_nodes and max_nodes > 0 else live


class DSElasticAgent(object):

    def __init__(self, cmd: List[str], env: Dict[str, str], node_id: int = 0, nproc_per_node: int = 1, min_nodes: int = 1, max_nodes: int = 1,
                 master_addr: str = "127.0.0.1", master_port: int = 29500, max_restarts: int = 100, monitor_interval: float = 1.0, store=None,
                 log_dir: Optional[str] = None):
        self.cmd = list(cmd)
        self.ds_env = dict(env)
        self.node_id = node_id
        self.nproc = nproc_per_node
        self.min_nodes, self.max_nodes = min_nodes, max_nodes
        self.master_addr, self.master_port = master_addr, master_port
        self.max_restarts = max_restarts
        self.remaining_restarts = max_restarts
        self.monitor_interval = monitor_interval
        self.store = store
        self.log_dir = log_dir
        self.procs: List[subprocess.Popen] = []
        self.epoch = -1
        self.members: List[int] = [node_id]

    # ------------------------------------------------------------------ membership over the store
    def _connect_store(self):
        if self.store is not None or self.max_nodes <= 1:
            return
        import torch.distributed as dist
        self.store = dist.TCPStore(self.master_addr, self.master_port + 1, is_master=(self.node_id == 0), timeout=timedelta(seconds=300),
                                   wait_for_workers=False)

    def _heartbeat(self):
        if self.store is not None:
            self.store.set(f"hb/{self.node_id}", str(time.time()))

    def _live_nodes(self):
        if self.store is None:
            return [self.node_id]
        live, now = [], time.time()
        for n in range(max(self.max_nodes, 1) * 4):
            try:
                if self.store.check([f"hb/{n}"]) and now - float(self.store.get(f"hb/{n}").decode()) < HEARTBEAT_TIMEOUT:
                    live.append(n)
            except Exception:
                pass
        return live

    def _sync_membership(self):
        """Returns True when a new epoch (different member list) was adopted."""
        if self.store is None:
            if self.epoch < 0:
                self.epoch, self.members = 0, [self.node_id]
                return True
            return False
        live = self._live_nodes()
        wanted = pick_members(live, self.min_nodes, self.max_nodes)
        if wanted is not None and self.node_id == min(live):
            current = self.store.get("members").decode().split(",") if self.store.check(["members"]) else []
            if [str(n) for n in wanted] != current:
                self.store.set("members", ",".join(str(n) for n in wanted))
                self.store.add("epoch", 1)
        if not self.store.check(["epoch"]):
            return False
        epoch = int(self.store.add("epoch", 0))
        if epoch != self.epoch:
            self.epoch = epoch
            self.members = [int(n) for n in self.store.get("members").decode().split(",") if n]
            return True
        return False

    # ------------------------------------------------------------------ workers
    def _worker_env(self, local_rank, global_rank, world_size, restart_count):
        env = copy.deepcopy(self.ds_env)
        env.update({
            "LOCAL_RANK": str(local_rank), "RANK": str(global_rank), "GROUP_RANK": str(sorted(self.members).index(self.node_id)),
            "ROLE_RANK": str(global_rank), "ROLE_NAME": "default", "LOCAL_WORLD_SIZE": str(self.nproc), "WORLD_SIZE": str(world_size),
            "GROUP_WORLD_SIZE": str(len(self.members)), "ROLE_WORLD_SIZE": str(world_size), "MASTER_ADDR": self.master_addr,
            "MASTER_PORT": str(self.master_port), "TORCHELASTIC_RESTART_COUNT": str(restart_count), "TORCHELASTIC_MAX_RESTARTS": str(self.max_restarts),
            "TORCHELASTIC_RUN_ID": env.get("TORCHELASTIC_RUN_ID", "ds_elastic"), "NCCL_ASYNC_ERROR_HANDLING": os.getenv("NCCL_ASYNC_ERROR_HANDLING", "1"),
        })
        if "OMP_NUM_THREADS" in os.environ:
            env["OMP_NUM_THREADS"] = os.environ["OMP_NUM_THREADS"]
        return env

    def _start_workers(self):
        restart_count = self.max_restarts - self.remaining_restarts
        if self.node_id not in self.members:
            logger.info(f"[elastic agent {self.node_id}] not a member of epoch {self.epoch}; standing by")
            self.procs = []
            return
        ranks, world = assign_ranks(self.members, self.node_id, self.nproc)
        self.procs = []
        for local_rank, rank in enumerate(ranks):
            cmd = [a.replace("${local_rank}", str(local_rank)) for a in self.cmd]
            out = None
            if self.log_dir:
                d = os.path.join(self.log_dir, f"attempt_{restart_count}")
                os.makedirs(d, exist_ok=True)
                out = open(os.path.join(d, f"rank{rank}.log"), "w")
            self.procs.append(subprocess.Popen(cmd, env=self._worker_env(local_rank, rank, world, restart_count), stdout=out, stderr=out))
        logger.info(f"[elastic agent {self.node_id}] epoch {self.epoch}: started ranks {ranks} of {world} (attempt {restart_count})")

    def _stop_workers(self):
        for p in self.procs:
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
        deadline = time.time() + 30
        for p in self.procs:
            try:
                p.wait(timeout=max(0.1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                p.kill()
        self.procs = []

    def _state(self):
        codes = [p.poll() for p in self.procs]
        if any(c is not None and c != 0 for c in codes):
            return "FAILED"
        if self.procs and all(c == 0 for c in codes):
            return "SUCCEEDED"
        return "HEALTHY"

    def run(self) -> int:
        self._connect_store()
        self._heartbeat()
        while not self._sync_membership():
            time.sleep(HEARTBEAT_INTERVAL / 4)
            self._heartbeat()
        self._start_workers()
        last_hb = time.time()
        while True:
            time.sleep(self.monitor_interval)
            if time.time() - last_hb > HEARTBEAT_INTERVAL:
                self._heartbeat()
                last_hb = time.time()
            state = self._state()
            if state == "SUCCEEDED":
                logger.info(f"[elastic agent {self.node_id}] worker group successfully finished.")
                return 0
            changed = self._sync_membership()
            if state == "FAILED" or changed:
                if state == "FAILED" and self.remaining_restarts <= 0:
                    self._stop_workers()
                    logger.error(f"[elastic agent {self.node_id}] worker group failed and no restarts remain")
                    return 1
                why = "membership changed" if changed else "worker failed"
                logger.info(f"[elastic agent {self.node_id}] {why}; {self.remaining_restarts}/{self.max_restarts} attempts left; restarting worker group")
                if state == "FAILED":
                    self.remaining_restarts -= 1
                self._stop_workers()
                self._start_workers()


def run_elastic(args, envs, current_env):
    """Entry used by launcher/launch.py --enable_elastic_training."""
    env = current_env.copy()
    env.update({k: v for k, v in envs[0].items() if k in ("CUDA_VISIBLE_DEVICES", "MASTER_ADDR", "MASTER_PORT")})
    cmd = []
    if not args.no_python:
        cmd += [sys.executable, "-u"] + (["-m"] if args.module else [])
    cmd += [args.training_script, "--local_rank=${local_rank}"] + list(args.training_script_args)
    agent = DSElasticAgent(cmd, env, node_id=args.node_rank, nproc_per_node=len(envs), min_nodes=max(1, args.min_elastic_nodes),
                           max_nodes=max(1, args.max_elastic_nodes), master_addr=args.master_addr, master_port=args.master_port)
    sys.exit(agent.run())
