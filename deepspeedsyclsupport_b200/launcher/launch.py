"""Per-node launcher (what `launcher/runner.py` or pdsh starts on every host).

    world info (urlsafe-base64 JSON {host: [gpu ids]})  ->  rank layout  ->  one child per local GPU  ->  supervise

* `build_rank_env` is the pure part: global ranks are numbered host by host in world-info order; each local process gets
  RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT plus CROSS_RANK (node index), CROSS_SIZE (node count) and
  LOCAL_SIZE, and CUDA_VISIBLE_DEVICES restricted to the node's selected GPUs.
* `NodeLauncher` builds each child's argv (optional numactl core binding, `python -u [-m] script --local_rank=N args...`),
  starts the children, and polls them once a second: the first non-zero exit tears the whole process tree down and becomes
  the launcher's exit code; SIGINT / SIGTERM do the same with exit code 1.
* Elastic jobs are handed to `elasticity.elastic_agent.run_elastic`.

Flags and environment names follow deepspeed/launcher/launch.py:130-355.
"""
import base64
import json
import os
import signal
import subprocess
import sys
import time
from argparse import REMAINDER, ArgumentParser
from collections import defaultdict

from ..utils.logging import logger

PID_FILE_BASEPATH = "/tmp"


def parse_args(args=None):
    ap = ArgumentParser(description="Start and supervise this node's share of a distributed job: one process per selected GPU")
    add = ap.add_argument
    add("--node_rank", type=int, default=0, help="index of this host in the world info")
    add("--master_addr", default="127.0.0.1", type=str)
    add("--master_port", default=29500, type=int)
    add("--world_info", default="None", type=str, help="urlsafe-base64 JSON {host: [gpu ids]}")
    add("--module", action="store_true")
    add("--no_python", action="store_true")
    add("--enable_elastic_training", action="store_true")
    add("--min_elastic_nodes", type=int, default=-1)
    add("--max_elastic_nodes", type=int, default=-1)
    add("--no_local_rank", action="store_true")
    add("--save_pid", type=int, default=0)
    add("--enable_each_rank_log", default="None", type=str)
    add("--bind_cores_to_rank", action="store_true")
    add("--bind_core_list", type=str, default=None)
    add("training_script", type=str)
    add("training_script_args", nargs=REMAINDER)
    return ap.parse_args(args=args)


def terminate_process_tree(pid):
    """SIGTERM `pid` and all its descendants, SIGKILL whatever is still alive after 30 s."""
    try:
        import psutil
    except ImportError:
        psutil = None
    if psutil is None:
        try:
            os.kill(pid, signal.SIGTERM)
        except ProcessLookupError:
            pass
        return
    try:
        root = psutil.Process(pid)
    except psutil.NoSuchProcess:
        return
    family = root.children(recursive=True) + [root]
    for proc in family:
        try:
            proc.terminate()
        except psutil.NoSuchProcess:
            pass
    _, survivors = psutil.wait_procs(family, timeout=30)
    for proc in survivors:
        try:
            proc.kill()
        except psutil.NoSuchProcess:
            pass


def build_rank_env(world_info, node_rank, master_addr, master_port):
    """-> ({host: [global ranks]}, [env dict of local process 0, 1, ...]) for host number `node_rank`."""
    hosts = list(world_info)
    mapping, next_rank = defaultdict(list), 0
    for host in hosts:
        count = len(world_info[host])
        mapping[host] = list(range(next_rank, next_rank + count))
        next_rank += count
    me = hosts[node_rank]
    shared = {"CUDA_VISIBLE_DEVICES": ",".join(str(g) for g in world_info[me]), "MASTER_ADDR": master_addr, "MASTER_PORT": str(master_port),
              "WORLD_SIZE": str(next_rank), "CROSS_RANK": str(node_rank), "CROSS_SIZE": str(len(hosts)), "LOCAL_SIZE": str(len(mapping[me]))}
    envs = [dict(shared, RANK=str(rank), LOCAL_RANK=str(local)) for local, rank in enumerate(mapping[me])]
    return mapping, envs


class NodeLauncher:
    """Children of one node: start, watch, stop."""

    def __init__(self, args, rank_envs, base_env):
        self.args, self.rank_envs, self.base_env = args, rank_envs, base_env
        self.children = []      # (Popen, argv)
        self.pid_file = None
        self.failed_with = None  # (argv, return code) of the first child that failed

    def child_command(self, local_rank):
        a = self.args
        if a.no_python and a.module:
            raise ValueError("Don't use both the '--no_python' flag and the '--module' flag at the same time.")
        extra_env, argv = {}, []
        if a.bind_cores_to_rank:
            from ..utils.numa import get_numactl_cmd
            cores, numactl = get_numactl_cmd(a.bind_core_list, len(self.rank_envs), local_rank)
            extra_env["OMP_NUM_THREADS"] = str(cores)
            argv += numactl
        if not a.no_python:
            argv += [sys.executable, "-u"] + (["-m"] if a.module else [])
        argv.append(a.training_script)
        if not a.no_local_rank:
            argv.append(f"--local_rank={local_rank}")
        return argv + list(a.training_script_args), extra_env

    def start(self):
        log_dir = None if self.args.enable_each_rank_log == "None" else self.args.enable_each_rank_log
        if log_dir:
            os.makedirs(log_dir, exist_ok=True)
        if self.args.save_pid:
            self.pid_file = os.path.join(PID_FILE_BASEPATH, f"{self.args.save_pid}.deepspeed")
            with open(self.pid_file, "w") as fd:
                fd.write(str(os.getpid()))
        for local_rank, rank_env in enumerate(self.rank_envs):
            argv, extra = self.child_command(local_rank)
            env = {**self.base_env, **rank_env, **extra}
            sink = open(os.path.join(log_dir, f"rank{rank_env['RANK']}.log"), "w") if log_dir else None
            self.children.append((subprocess.Popen(argv, env=env, stdout=sink, stderr=sink), argv))

    def stop(self, signum=None, frame=None):
        """Kill every child tree and exit: with the failed child's code if one failed, else 1."""
        for proc, _ in self.children:
            logger.info(f"Killing subprocess {proc.pid}")
            try:
                terminate_process_tree(proc.pid)
            except Exception:
                pass
        if self.failed_with is not None:
            argv, code = self.failed_with
            logger.error(f"{argv} exits with return code = {code}")
            sys.exit(code)
        if signum is not None:
            logger.info(f"Main process received {signal.Signals(signum).name}, exiting")
        if self.pid_file and os.path.isfile(self.pid_file):
            os.remove(self.pid_file)
        sys.exit(1)

    def supervise(self):
        for s in (signal.SIGINT, signal.SIGTERM):
            signal.signal(s, self.stop)
        running = list(self.children)
        while running:
            still = []
            for proc, argv in running:
                code = proc.poll()
                if code is None:
                    still.append((proc, argv))
                elif code != 0:
                    self.failed_with = (argv, code)
                    self.stop()
                else:
                    logger.info(f"Process {proc.pid} exits successfully.")
            running = still
            if running:
                time.sleep(1)


def main():
    args = parse_args()
    base_env = os.environ.copy()
    for name, value in base_env.items():
        if "NCCL" in name:
            logger.info(f"{args.node_rank} {name}={value}")
    if args.world_info == "None":
        raise ValueError("world_info can not be None")
    world_info = json.loads(base64.urlsafe_b64decode(args.world_info))
    logger.info(f"WORLD INFO DICT: {world_info}")
    mapping, rank_envs = build_rank_env(world_info, args.node_rank, args.master_addr, args.master_port)
    logger.info(f"global_rank_mapping={dict(mapping)}")
    if args.enable_elastic_training:
        from ..elasticity.elastic_agent import run_elastic
        return run_elastic(args, rank_envs, base_env)
    node = NodeLauncher(args, rank_envs, base_env)
    node.start()
    node.supervise()


if __name__ == "__main__":
    main()
