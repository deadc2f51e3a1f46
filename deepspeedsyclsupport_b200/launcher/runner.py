"""`deepspeed` launcher CLI: hostfile / --include / --exclude parsing -> base64 world info -> one `launch.py` per node
(directly on a single node, through PDSH / OpenMPI / MPICH / IMPI / Slurm / MVAPICH across nodes).
Parity: deepspeed/launcher/runner.py:45-597."""
import argparse
import base64
import collections
import json
import os
import re
import shutil
import signal
import subprocess
import sys
import time
from copy import deepcopy

from ..accelerator import get_accelerator
from ..utils.logging import logger
from .constants import (ELASTIC_TRAINING_ID_DEFAULT, IMPI_LAUNCHER, MPICH_LAUNCHER, MVAPICH_LAUNCHER, OPENMPI_LAUNCHER, PDSH_LAUNCHER,
                        SLURM_LAUNCHER)
from .multinode_runner import IMPIRunner, MPICHRunner, MVAPICHRunner, OpenMPIRunner, PDSHRunner, SlurmRunner

DLTS_HOSTFILE = "/job/hostfile"
EXPORT_ENVS = ["MLFLOW", "NCCL", "PYTHON", "MV2", "UCX", "CUDA", "DSB200"]
DEEPSPEED_ENVIRONMENT_NAME = os.getenv("DS_ENV_FILE", ".deepspeed_env")
DEEPSPEED_ENVIRONMENT_PATHS = [os.path.expanduser("~"), "."]
PDSH_MAX_FAN_OUT = 1024
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500


def parse_args(args=None):
    parser = argparse.ArgumentParser(description="DeepSpeed runner to help launch distributed multi-node/multi-gpu training jobs.")
    parser.add_argument("-H", "--hostfile", type=str, default=DLTS_HOSTFILE,
                        help="Hostfile path (in MPI style) that defines the resource pool available to the job (e.g., worker-0 slots=4)")
    parser.add_argument("-i", "--include", type=str, default="",
                        help="Specify hardware resources to use during execution. NODE_SPEC[@NODE_SPEC ...] where NODE_SPEC=NAME[:SLOT[,SLOT ...]].")
    parser.add_argument("-e", "--exclude", type=str, default="", help="Specify hardware resources to NOT use during execution. Mutually exclusive with --include.")
    parser.add_argument("--num_nodes", type=int, default=-1, help="Total number of worker nodes to run on, this will use the top N hosts from the given hostfile.")
    parser.add_argument("--min_elastic_nodes", type=int, default=-1)
    parser.add_argument("--max_elastic_nodes", type=int, default=-1)
    parser.add_argument("--num_gpus", "--num_accelerators", type=int, default=-1, help="Max number of GPUs to use on each node, will use [0:N) GPU ids on each node.")
    parser.add_argument("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int)
    parser.add_argument("--master_addr", default="", type=str)
    parser.add_argument("--launcher", default=PDSH_LAUNCHER, type=str,
                        help="(optional) choose launcher backend for multi-node training. Options currently include PDSH, OpenMPI, MVAPICH, SLURM, MPICH, IMPI.")
    parser.add_argument("--launcher_args", default="", type=str)
    parser.add_argument("--module", action="store_true", help="Change each process to interpret the launch script as a Python module, executing with the same behavior as 'python -m'.")
    parser.add_argument("--no_python", action="store_true", help="Skip prepending the training script with 'python' - just execute it directly.")
    parser.add_argument("--no_local_rank", action="store_true", help="Do not pass local_rank as an argument when calling the user's training script.")
    parser.add_argument("--no_ssh_check", action="store_true")
    parser.add_argument("--force_multi", action="store_true", help="Force multi-node launcher mode, helps in cases where user wants to launch on single remote node.")
    parser.add_argument("--save_pid", action="store_true")
    parser.add_argument("--enable_each_rank_log", default="None", type=str)
    parser.add_argument("--autotuning", default="", choices=["tune", "run"], type=str)
    parser.add_argument("--elastic_training", action="store_true")
    parser.add_argument("--bind_cores_to_rank", action="store_true")
    parser.add_argument("--bind_core_list", type=str, default=None)
    parser.add_argument("--ssh_port", type=int, default=None)
    parser.add_argument("user_script", type=str, help="User script to launch, followed by any required arguments.")
    parser.add_argument("user_args", nargs=argparse.REMAINDER)
    return parser.parse_args(args=args)


def fetch_hostfile(hostfile_path):
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    with open(hostfile_path, "r") as fd:
        hostfile_text = fd.readlines()
    return _parse_hostfile(hostfile_text)


def _parse_hostfile(hostfile_lines):
    pattern = r"^(\S+)\s+slots=(\d+)"
    resource_pool = collections.OrderedDict()
    for line in hostfile_lines:
        line = line.strip()
        match = re.search(pattern, line)
        if line.startswith("#") or line == "":
            continue
        elif match:
            host = match.group(1)
            num_slots = int(match.group(2))
            if host in resource_pool:
                logger.error(f"Bad hostfile text: {hostfile_lines}")
                raise ValueError(f"Hostfile contains multiple entries for {host}, unable to proceed with launching")
            resource_pool[host] = num_slots
        else:
            logger.error(f"Bad hostfile text: {hostfile_lines}")
            raise ValueError(f"Hostfile contains a bad entry: {line}, unable to proceed with launching")
    if len(resource_pool) == 0:
        logger.error(f"Bad hostfile text: {hostfile_lines}")
        raise ValueError("Hostfile is empty or not formatted correctly, unable to proceed with launching.")
    return resource_pool


def _stable_remove_duplicates(data):
    new_list = []
    for x in data:
        if x not in new_list:
            new_list.append(x)
    return new_list


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """Filter {host: [slots]} with an include or exclude string: 'worker-0@worker-1:0,2' style."""
    NODE_SEP = "@"
    SLOT_LIST_START = ":"
    SLOT_SEP = ","
    if include_str == "" and exclude_str == "":
        return host_info
    if include_str != "" and exclude_str != "":
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    filtered_hosts = dict()
    if include_str:
        parse_str = include_str
    if exclude_str != "":
        filtered_hosts = deepcopy(host_info)
        parse_str = exclude_str
    for node_config in parse_str.split(NODE_SEP):
        if SLOT_LIST_START in node_config:
            hostname, slots = node_config.split(SLOT_LIST_START)
            slots = [int(x) for x in slots.split(SLOT_SEP)]
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            for slot in slots:
                if slot not in host_info[hostname]:
                    raise ValueError(f"No slot '{slot}' specified on host '{hostname}'")
            if include_str:
                filtered_hosts[hostname] = slots
            elif exclude_str:
                for slot in slots:
                    logger.info(f"removing {slot} from {hostname}")
                    filtered_hosts[hostname].remove(slot)
        else:
            hostname = node_config
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            if include_str:
                filtered_hosts[hostname] = host_info[hostname]
            elif exclude_str:
                filtered_hosts[hostname] = []
    del_keys = []
    for hostname in filtered_hosts:
        filtered_hosts[hostname] = _stable_remove_duplicates(filtered_hosts[hostname])
        if len(filtered_hosts[hostname]) == 0:
            del_keys.append(hostname)
    for name in del_keys:
        del filtered_hosts[name]
    ordered_hosts = collections.OrderedDict()
    for host in host_info:
        if host in filtered_hosts:
            ordered_hosts[host] = filtered_hosts[host]
    return ordered_hosts


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    active_resources = collections.OrderedDict()
    for hostname, slots in resource_pool.items():
        active_resources[hostname] = list(range(slots))
    return parse_resource_filter(active_resources, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info):
    world_info_json = json.dumps(world_info).encode("utf-8")
    return base64.urlsafe_b64encode(world_info_json).decode("utf-8")


def run_autotuning(args, active_resources):
    from ..autotuning.autotuner import Autotuner
    tuner = Autotuner(args, active_resources)
    logger.info("[Start] Running autotuning")
    tuner.tune()
    tuner.print_tuning_results()
    logger.info("[End] Running autotuning")
    tuner.write_optimal_config()
    if args.autotuning == "run":
        tuner.run_after_tuning()


def parse_num_nodes(str_num_nodes: str, elastic_training: bool):
    node_list = str_num_nodes.split(":")
    if len(node_list) == 1:
        min_nodes, max_nodes = int(node_list[0]), -1
    elif len(node_list) == 2 and elastic_training:
        min_nodes, max_nodes = int(node_list[0]), int(node_list[1])
    elif len(node_list) == 2 and not elastic_training:
        raise RuntimeError("MIN:MAX format is only supported in elastic training")
    else:
        raise RuntimeError("num_nodes {} is not in MIN:MAX format".format(str_num_nodes))
    return min_nodes, max_nodes


def build_launch_command(args, active_resources, world_info_base64):
    """The per-node launch command for the single-node path (separated so it can be unit-tested)."""
    cmd = [sys.executable, "-u", "-m", "deepspeedsyclsupport_b200.launcher.launch", f"--world_info={world_info_base64}",
           f"--master_addr={args.master_addr}", f"--master_port={args.master_port}"]
    if args.no_python:
        cmd += ["--no_python"]
    if args.no_local_rank:
        cmd += ["--no_local_rank"]
    if args.module:
        cmd += ["--module"]
    if args.save_pid:
        cmd += ["--save_pid", f"{os.getpid()}"]
    if args.enable_each_rank_log != "None":
        cmd.append(f"--enable_each_rank_log={args.enable_each_rank_log}")
    if args.elastic_training:
        cmd.append("--enable_elastic_training")
        cmd.append(f"--max_elastic_nodes={args.max_elastic_nodes}")
        cmd.append(f"--min_elastic_nodes={args.min_elastic_nodes}")
    if args.bind_cores_to_rank:
        cmd.append("--bind_cores_to_rank")
    if args.bind_core_list is not None:
        cmd.append(f"--bind_core_list={args.bind_core_list}")
    return cmd + [args.user_script] + args.user_args


def main(args=None):
    args = parse_args(args)
    if args.elastic_training:
        assert args.master_addr != "", "Master Addr is required when elastic training is enabled"
    resource_pool = fetch_hostfile(args.hostfile)
    if not resource_pool and len(os.environ.get("CUDA_VISIBLE_DEVICES", "")):
        raise ValueError("[Error] CUDA_VISIBLE_DEVICES is set: use the --include localhost:... syntax to select devices")
    multi_node_exec = True
    if not resource_pool:
        resource_pool = {}
        device_count = get_accelerator().device_count()
        if device_count == 0:
            raise RuntimeError("Unable to proceed, no GPU resources available")
        resource_pool["localhost"] = device_count
        args.master_addr = "127.0.0.1"
        multi_node_exec = False
    if not multi_node_exec and args.num_nodes > 1:
        raise ValueError("Num nodes is >1 but no extra nodes available via hostfile")
    active_resources = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    env = os.environ.copy()
    if not args.master_addr:
        assert multi_node_exec
        first_host = list(active_resources.keys())[0]
        try:
            hostname_cmd = ["ssh", first_host, "hostname -I"] if args.ssh_port is None else ["ssh", "-p", str(args.ssh_port), first_host, "hostname -I"]
            result = subprocess.check_output(hostname_cmd)
        except subprocess.CalledProcessError as err:
            logger.error("Unable to detect suitable master address via `hostname -I`, please manually specify one via --master_addr")
            raise err
        args.master_addr = result.decode("utf-8").split()[0]
        if not args.master_addr:
            raise RuntimeError("Unable to detect suitable master address via `hostname -I`, please manually specify one via --master_addr")
        logger.info(f"Using IP address of {args.master_addr} for node {first_host}")
    if args.autotuning != "":
        run_autotuning(args, active_resources)
        return
    if args.num_nodes > 0:
        updated_active_resources = collections.OrderedDict()
        for count, hostname in enumerate(active_resources.keys()):
            if args.num_nodes == count:
                break
            updated_active_resources[hostname] = active_resources[hostname]
        active_resources = updated_active_resources
    if args.num_gpus > 0:
        updated_active_resources = collections.OrderedDict()
        for hostname in active_resources.keys():
            updated_active_resources[hostname] = list(range(args.num_gpus))
        active_resources = updated_active_resources
    if args.elastic_training:
        assert not args.no_local_rank, "--no_local_rank argument is not supported in Elastic training"
    world_info_base64 = encode_world_info(active_resources)
    multi_node_exec = args.force_multi or len(active_resources) > 1
    if not multi_node_exec:
        cmd = build_launch_command(args, active_resources, world_info_base64)
    else:
        args.launcher = args.launcher.lower()
        runner_cls = {PDSH_LAUNCHER: PDSHRunner, OPENMPI_LAUNCHER: OpenMPIRunner, MPICH_LAUNCHER: MPICHRunner, IMPI_LAUNCHER: IMPIRunner,
                      MVAPICH_LAUNCHER: MVAPICHRunner, SLURM_LAUNCHER: SlurmRunner}.get(args.launcher)
        if runner_cls is None:
            raise NotImplementedError(f"Unknown launcher {args.launcher}")
        runner = runner_cls(args, world_info_base64, resource_pool)
        if not runner.backend_exists():
            raise RuntimeError(f"launcher '{args.launcher}' not installed.")
        curr_path = os.path.abspath(".")
        env["PYTHONPATH"] = curr_path + ":" + env["PYTHONPATH"] if "PYTHONPATH" in env else curr_path
        excluded_vars = []
        for exclude_key, var_list in {"EXCLUDE_ENVS": ["PYTHONPATH", "PATH"]}.items():
            excluded_vars += var_list
        exports = EXPORT_ENVS + get_accelerator().export_envs()
        for var in env.keys():
            if any([var.startswith(name) for name in exports]):
                if not any([var == name for name in excluded_vars]):
                    runner.add_export(var, env[var])
        for environ_path in DEEPSPEED_ENVIRONMENT_PATHS:
            environ_file = os.path.join(environ_path, DEEPSPEED_ENVIRONMENT_NAME)
            if os.path.isfile(environ_file):
                logger.info(f"deepspeed_env file = {environ_file}")
                with open(environ_file, "r") as fd:
                    for var in fd.readlines():
                        key, val = var.split("=", maxsplit=1)
                        runner.add_export(key, val)
        cmd, kill_cmd, env = runner.get_cmd(env, active_resources)
    logger.info(f"cmd = {' '.join(cmd)}")
    result = subprocess.Popen(cmd, env=env)

    def sigkill_handler(signum, frame):
        result.send_signal(signal.SIGINT)
        time.sleep(0.1)
        result.send_signal(signal.SIGTERM)
        result_kill = subprocess.Popen(kill_cmd, env=env) if multi_node_exec and "kill_cmd" in dir() else None
        if result_kill:
            result_kill.wait()
        time.sleep(1)
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
