"""Per-node launcher: decodes the world info, exports RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (+ CROSS_RANK,
CROSS_SIZE, LOCAL_SIZE), spawns one process per local GPU (optionally numactl-bound), polls the children and kills the
whole tree as soon as one fails.  Elastic mode hands the job to DSElasticAgent.
Parity: deepspeed/launcher/launch.py:130-355."""
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
    parser = ArgumentParser(description="DeepSpeed distributed training launch utility that creates multiple distributed processes on a single node")
    parser.add_argument("--node_rank", type=int, default=0, help="The rank of the node for multi-node distributed training")
    parser.add_argument("--master_addr", default="127.0.0.1", type=str)
    parser.add_argument("--master_port", default=29500, type=int)
    parser.add_argument("--world_info", default="None", type=str, help="world info base64 encoded dictionary")
    parser.add_argument("--module", action="store_true")
    parser.add_argument("--no_python", action="store_true")
    parser.add_argument("--enable_elastic_training", action="store_true")
    parser.add_argument("--min_elastic_nodes", type=int, default=-1)
    parser.add_argument("--max_elastic_nodes", type=int, default=-1)
    parser.add_argument("--no_local_rank", action="store_true")
    parser.add_argument("--save_pid", type=int, default=0)
    parser.add_argument("--enable_each_rank_log", default="None", type=str)
    parser.add_argument("--bind_cores_to_rank", action="store_true")
    parser.add_argument("--bind_core_list", type=str, default=None)
    parser.add_argument("training_script", type=str)
    parser.add_argument("training_script_args", nargs=REMAINDER)
    return parser.parse_args(args=args)


def terminate_process_tree(pid):
    try:
        import psutil
        process = psutil.Process(pid)
        children = process.children(recursive=True)
        children.append(process)
        for child in children:
            try:
                child.terminate()
            except psutil.NoSuchProcess:
                pass
        gone, alive = psutil.wait_procs(children, timeout=30)
        for p in alive:
            p.kill()
    except Exception:
        try:
            os.kill(pid, signal.SIGTERM)
        except ProcessLookupError:
            pass


def build_rank_env(world_info, node_rank, master_addr, master_port):
    """(global_rank_mapping, per-local-rank env dicts) for this node -- pure function for tests."""
    node_list = list(world_info.keys())
    local_node = node_list[node_rank]
    local_gpu_ids = world_info[local_node]
    num_local_procs = len(local_gpu_ids)
    global_rank_mapping = defaultdict(list)
    curr_global_rank = 0
    dist_world_size = 0
    for node_id in node_list:
        gids = world_info[node_id]
        dist_world_size += len(gids)
        for gid in gids:
            global_rank_mapping[node_id].append(curr_global_rank)
            curr_global_rank += 1
    envs = []
    for local_proc in range(num_local_procs):
        e = {"CUDA_VISIBLE_DEVICES": ",".join(map(str, local_gpu_ids)), "MASTER_ADDR": master_addr, "MASTER_PORT": str(master_port),
             "WORLD_SIZE": str(dist_world_size), "CROSS_RANK": str(node_rank), "CROSS_SIZE": str(len(node_list)),
             "LOCAL_SIZE": str(num_local_procs), "RANK": str(global_rank_mapping[local_node][local_proc]), "LOCAL_RANK": str(local_proc)}
        envs.append(e)
    return global_rank_mapping, envs


def main():
    args = parse_args()
    current_env = os.environ.copy()
    for k in current_env.keys():
        if "NCCL" in k:
            logger.info(f"{args.node_rank} {k}={current_env[k]}")
    if args.world_info == "None":
        raise ValueError("world_info can not be None")
    world_info = json.loads(base64.urlsafe_b64decode(args.world_info))
    logger.info(f"WORLD INFO DICT: {world_info}")
    global_rank_mapping, envs = build_rank_env(world_info, args.node_rank, args.master_addr, args.master_port)
    logger.info(f"global_rank_mapping={dict(global_rank_mapping)}")
    if args.save_pid:
        launcher_pid = os.getpid()
        pid_file = os.path.join(PID_FILE_BASEPATH, f"{args.save_pid}.deepspeed")
        with open(pid_file, "w") as fd:
            fd.write(f"{launcher_pid}")
    processes = []
    cmd = []
    if not args.enable_elastic_training:
        if args.enable_each_rank_log != "None":
            os.makedirs(args.enable_each_rank_log, exist_ok=True)
        for local_proc, e in enumerate(envs):
            env = current_env.copy()
            env.update(e)
            cmd = []
            if args.bind_cores_to_rank:
                from ..utils.numa import get_numactl_cmd
                cores_per_rank, numactl_cmd = get_numactl_cmd(args.bind_core_list, len(envs), local_proc)
                env["OMP_NUM_THREADS"] = f"{cores_per_rank}"
                cmd = cmd + numactl_cmd
            if not args.no_python:
                cmd.append(sys.executable)
                cmd.append("-u")
                if args.module:
                    cmd.append("-m")
            else:
                if args.module:
                    raise ValueError("Don't use both the '--no_python' flag and the '--module' flag at the same time.")
            cmd.append(args.training_script)
            if not args.no_local_rank:
                cmd.append(f"--local_rank={local_proc}")
            cmd += args.training_script_args
            if args.enable_each_rank_log != "None":
                log_file = os.path.join(args.enable_each_rank_log, f"rank{e['RANK']}.log")
                log_fd = open(log_file, "w")
                process = subprocess.Popen(cmd, env=env, stdout=log_fd, stderr=log_fd)
            else:
                process = subprocess.Popen(cmd, env=env)
            processes.append(process)
    else:
        from ..elasticity import DSElasticAgent
        from ..elasticity.elastic_agent import run_elastic
        run_elastic(args, envs, current_env)
        return

    sig_names = {2: "SIGINT", 15: "SIGTERM"}
    last_return_code = None

    def sigkill_handler(signum, frame):
        for process in processes:
            logger.info(f"Killing subprocess {process.pid}")
            try:
                terminate_process_tree(process.pid)
            except Exception:
                pass
        if last_return_code is not None:
            logger.error(f"{cmd} exits with return code = {last_return_code}")
            sys.exit(last_return_code)
        if signum in sig_names:
            logger.info(f"Main process received {sig_names[signum]}, exiting")
        if args.save_pid:
            if os.path.isfile(pid_file):
                os.remove(pid_file)
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    alive_processes = set(processes)
    while len(alive_processes):
        finished_processes = []
        for process in alive_processes:
            if process.poll() is None:
                continue
            if process.returncode != 0:
                last_return_code = process.returncode
                sigkill_handler(signal.SIGTERM, None)
            else:
                finished_processes.append(process)
                logger.info(f"Process {process.pid} exits successfully.")
        alive_processes = set(alive_processes) - set(finished_processes)
        time.sleep(1)


if __name__ == "__main__":
    main()
