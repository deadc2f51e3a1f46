"""`deepspeed` command line: decide WHICH slots of WHICH hosts take part, then start one `launcher.launch` per node.

Stages of `main()`:
    1. resources   hostfile ("host slots=N" lines) or, without one, the local GPUs
    2. selection   --include / --exclude (NODE[:SLOT,...]@NODE...), --num_nodes, --num_gpus  -> {host: [slot ids]}
    3. rendezvous  master address (given, 127.0.0.1 locally, or `hostname -I` of the first host over ssh)
    4. command     single node: exec launch.py here; several nodes: a MultiNodeRunner (pdsh / MPI flavours / slurm) builds
                   one command that fans out, with the whitelisted environment exported
    5. supervise   forward SIGINT/SIGTERM to the child (and run the runner's kill command), propagate the exit code

The world description handed to every node is urlsafe-base64 JSON of the selection (`encode_world_info`), the format
`launcher/launch.py` and the reference's launcher agree on.  CLI flags follow deepspeed/launcher/runner.py:45-597 so existing
job scripts keep working.
"""
import argparse
import base64
import json
import os
import re
import signal
import subprocess
import sys
import time
from collections import OrderedDict

from ..accelerator import get_accelerator
from ..utils.logging import logger
from .constants import (ELASTIC_TRAINING_ID_DEFAULT, IMPI_LAUNCHER, MPICH_LAUNCHER, MVAPICH_LAUNCHER, OPENMPI_LAUNCHER, PDSH_LAUNCHER,
                        SLURM_LAUNCHER)
from .multinode_runner import IMPIRunner, MPICHRunner, MVAPICHRunner, OpenMPIRunner, PDSHRunner, SlurmRunner

DLTS_HOSTFILE = "/job/hostfile"
# environment variables whose NAME starts with one of these travel to the remote nodes (plus the accelerator's own list)
EXPORT_ENVS = ["MLFLOW", "NCCL", "PYTHON", "MV2", "UCX", "CUDA", "DSB200"]
NEVER_EXPORTED = ("PYTHONPATH", "PATH")
DEEPSPEED_ENVIRONMENT_NAME = os.getenv("DS_ENV_FILE", ".deepspeed_env")
DEEPSPEED_ENVIRONMENT_PATHS = [os.path.expanduser("~"), "."]
PDSH_MAX_FAN_OUT = 1024
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500

_RUNNERS = {PDSH_LAUNCHER: PDSHRunner, OPENMPI_LAUNCHER: OpenMPIRunner, MPICH_LAUNCHER: MPICHRunner, IMPI_LAUNCHER: IMPIRunner,
            MVAPICH_LAUNCHER: MVAPICHRunner, SLURM_LAUNCHER: SlurmRunner}


# ------------------------------------------------------------------ command line
def parse_args(args=None):
    ap = argparse.ArgumentParser(description="Start a (multi-node) multi-GPU job: one process per GPU slot on every selected host.")
    add = ap.add_argument
    add("-H", "--hostfile", type=str, default=DLTS_HOSTFILE, help="MPI-style resource file, one 'hostname slots=N' line per node")
    add("-i", "--include", type=str, default="", help="use only these resources: NODE[:SLOT[,SLOT...]][@NODE...]")
    add("-e", "--exclude", type=str, default="", help="use everything but these resources (same syntax; not together with --include)")
    add("--num_nodes", type=int, default=-1, help="keep the first N hosts of the selection")
    add("--min_elastic_nodes", type=int, default=-1)
    add("--max_elastic_nodes", type=int, default=-1)
    add("--num_gpus", "--num_accelerators", type=int, default=-1, help="use slots [0, N) on every selected host")
    add("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int)
    add("--master_addr", default="", type=str)
    add("--launcher", default=PDSH_LAUNCHER, type=str, help="multi-node transport: " + ", ".join(sorted(_RUNNERS)))
    add("--launcher_args", default="", type=str)
    add("--module", action="store_true", help="run the script as `python -m <script>`")
    add("--no_python", action="store_true", help="execute the script directly instead of through the interpreter")
    add("--no_local_rank", action="store_true", help="do not append --local_rank=<n> to the script's arguments")
    add("--no_ssh_check", action="store_true")
    add("--force_multi", action="store_true", help="take the multi-node path even for a single (remote) host")
    add("--save_pid", action="store_true")
    add("--enable_each_rank_log", default="None", type=str)
    add("--autotuning", default="", choices=["tune", "run"], type=str)
    add("--elastic_training", action="store_true")
    add("--bind_cores_to_rank", action="store_true")
    add("--bind_core_list", type=str, default=None)
    add("--ssh_port", type=int, default=None)
    add("user_script", type=str, help="the training script, followed by its own arguments")
    add("user_args", nargs=argparse.REMAINDER)
    return ap.parse_args(args=args)


# ------------------------------------------------------------------ stage 1: resources
_HOST_LINE = re.compile(r"^(\S+)\s+slots=(\d+)")


def _parse_hostfile(hostfile_lines):
    """['host slots=N', ...] -> OrderedDict host -> N.  Blank lines and '#' comments are skipped; anything else that does
    not parse, a repeated host, or an empty result is an error."""
    pool = OrderedDict()

    def bad(why):
        logger.error(f"Bad hostfile text: {hostfile_lines}")
        return ValueError(f"Hostfile {why}, unable to proceed with launching")

    for raw in hostfile_lines:
        text = raw.strip()
        if not text or text.startswith("#"):
            continue
        m = _HOST_LINE.search(text)
        if m is None:
            raise bad(f"contains a bad entry: {text}")
        host, slots = m.group(1), int(m.group(2))
        if host in pool:
            raise bad(f"contains multiple entries for {host}")
        pool[host] = slots
    if not pool:
        raise bad("is empty or not formatted correctly")
    return pool


def fetch_hostfile(hostfile_path):
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    with open(hostfile_path) as fd:
        return _parse_hostfile(fd.readlines())


# ------------------------------------------------------------------ stage 2: selection
def _parse_node_specs(spec, host_info):
    """'a@b:0,2' -> [('a', None), ('b', [0, 2])]; every host / slot must exist in `host_info`."""
    out = []
    for item in spec.split("@"):
        host, _, slot_text = item.partition(":")
        if host not in host_info:
            raise ValueError(f"Hostname '{host}' not found in hostfile")
        if not slot_text:
            out.append((host, None))
            continue
        slots = [int(s) for s in slot_text.split(",")]
        for s in slots:
            if s not in host_info[host]:
                raise ValueError(f"No slot '{s}' specified on host '{host}'")
        out.append((host, slots))
    return out


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """Apply an include OR an exclude string to {host: [slot ids]}.  Host order of `host_info` is kept, hosts left without
    slots disappear, repeated slots collapse (first occurrence wins)."""
    if include_str and exclude_str:
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    if not include_str and not exclude_str:
        return host_info
    if include_str:
        chosen = {}
        for host, slots in _parse_node_specs(include_str, host_info):
            chosen[host] = list(host_info[host]) if slots is None else slots
    else:
        chosen = {h: list(s) for h, s in host_info.items()}
        for host, slots in _parse_node_specs(exclude_str, host_info):
            if slots is None:
                chosen[host] = []
                continue
            for s in slots:
                logger.info(f"removing {s} from {host}")
                chosen[host].remove(s)
    result = OrderedDict()
    for host in host_info:
        slots = list(OrderedDict.fromkeys(chosen.get(host, ())))
        if slots:
            result[host] = slots
    return result


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    every_slot = OrderedDict((host, list(range(n))) for host, n in resource_pool.items())
    return parse_resource_filter(every_slot, include_str=inclusion, exclude_str=exclusion)


def _limit_selection(active, num_nodes, num_gpus):
    if num_nodes > 0:
        active = OrderedDict(list(active.items())[:num_nodes])
    if num_gpus > 0:
        active = OrderedDict((host, list(range(num_gpus))) for host in active)
    return active


def parse_num_nodes(str_num_nodes: str, elastic_training: bool):
    """'N' -> (N, -1); 'MIN:MAX' -> (MIN, MAX), elastic jobs only."""
    parts = str_num_nodes.split(":")
    if len(parts) > 2:
        raise RuntimeError("num_nodes {} is not in MIN:MAX format".format(str_num_nodes))
    if len(parts) == 2 and not elastic_training:
        raise RuntimeError("MIN:MAX format is only supported in elastic training")
    return int(parts[0]), (int(parts[1]) if len(parts) == 2 else -1)


def encode_world_info(world_info):
    return base64.urlsafe_b64encode(json.dumps(world_info).encode("utf-8")).decode("utf-8")


# ------------------------------------------------------------------ stage 3: rendezvous address
def _first_host_address(host, ssh_port):
    ssh = ["ssh"] + (["-p", str(ssh_port)] if ssh_port is not None else []) + [host, "hostname -I"]
    hint = "Unable to detect suitable master address via `hostname -I`, please manually specify one via --master_addr"
    try:
        words = subprocess.check_output(ssh).decode("utf-8").split()
    except subprocess.CalledProcessError:
        logger.error(hint)
        raise
    if not words:
        raise RuntimeError(hint)
    logger.info(f"Using IP address of {words[0]} for node {host}")
    return words[0]


# ------------------------------------------------------------------ stage 4: commands
def build_launch_command(args, active_resources, world_info_base64):
    """argv that starts the ranks of THIS node (single-node path; also what the tests inspect)."""
    switches = [("--no_python", args.no_python), ("--no_local_rank", args.no_local_rank), ("--module", args.module),
                ("--bind_cores_to_rank", args.bind_cores_to_rank)]
    cmd = [sys.executable, "-u", "-m", "deepspeedsyclsupport_b200.launcher.launch", f"--world_info={world_info_base64}",
           f"--master_addr={args.master_addr}", f"--master_port={args.master_port}"]
    cmd += [flag for flag, on in switches[:3] if on]
    if args.save_pid:
        cmd += ["--save_pid", str(os.getpid())]
    if args.enable_each_rank_log != "None":
        cmd += [f"--enable_each_rank_log={args.enable_each_rank_log}"]
    if args.elastic_training:
        cmd += ["--enable_elastic_training", f"--max_elastic_nodes={args.max_elastic_nodes}", f"--min_elastic_nodes={args.min_elastic_nodes}"]
    cmd += [flag for flag, on in switches[3:] if on]
    if args.bind_core_list is not None:
        cmd += [f"--bind_core_list={args.bind_core_list}"]
    return cmd + [args.user_script] + list(args.user_args)


def _exported_environment(env):
    """(name, value) pairs that must exist on the remote side: whitelisted prefixes, then the lines of a .deepspeed_env file."""
    prefixes = tuple(EXPORT_ENVS + list(get_accelerator().export_envs()))
    for name, value in env.items():
        if name.startswith(prefixes) and name not in NEVER_EXPORTED:
            yield name, value
    for folder in DEEPSPEED_ENVIRONMENT_PATHS:
        path = os.path.join(folder, DEEPSPEED_ENVIRONMENT_NAME)
        if os.path.isfile(path):
            logger.info(f"deepspeed_env file = {path}")
            with open(path) as fd:
                for line in fd:
                    if "=" in line:
                        name, value = line.split("=", maxsplit=1)
                        yield name, value


def _multi_node_command(args, env, resource_pool, active, world_info_base64):
    args.launcher = args.launcher.lower()
    if args.launcher not in _RUNNERS:
        raise NotImplementedError(f"Unknown launcher {args.launcher}")
    runner = _RUNNERS[args.launcher](args, world_info_base64, resource_pool)
    if not runner.backend_exists():
        raise RuntimeError(f"launcher '{args.launcher}' not installed.")
    here = os.path.abspath(".")
    env["PYTHONPATH"] = here + (":" + env["PYTHONPATH"] if "PYTHONPATH" in env else "")
    for name, value in _exported_environment(env):
        runner.add_export(name, value)
    return runner.get_cmd(env, active)


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


# ------------------------------------------------------------------ stage 5: supervise
def _supervise(cmd, env, kill_cmd):
    child = subprocess.Popen(cmd, env=env)

    def stop(signum, frame):
        child.send_signal(signal.SIGINT)
        time.sleep(0.1)
        child.send_signal(signal.SIGTERM)
        if kill_cmd:
            subprocess.Popen(kill_cmd, env=env).wait()
        time.sleep(1)
        sys.exit(1)

    for s in (signal.SIGINT, signal.SIGTERM):
        signal.signal(s, stop)
    child.wait()
    return child.returncode


def main(args=None):
    args = parse_args(args)
    if args.elastic_training:
        assert args.master_addr != "", "Master Addr is required when elastic training is enabled"
        assert not args.no_local_rank, "--no_local_rank argument is not supported in Elastic training"
    # 1. resources
    resource_pool = fetch_hostfile(args.hostfile)
    local_only = not resource_pool
    if local_only:
        if os.environ.get("CUDA_VISIBLE_DEVICES", ""):
            raise ValueError("[Error] CUDA_VISIBLE_DEVICES is set: use the --include localhost:... syntax to select devices")
        n = get_accelerator().device_count()
        if n == 0:
            raise RuntimeError("Unable to proceed, no GPU resources available")
        if args.num_nodes > 1:
            raise ValueError("Num nodes is >1 but no extra nodes available via hostfile")
        resource_pool = {"localhost": n}
        args.master_addr = "127.0.0.1"
    # 2. selection
    active = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    # 3. rendezvous
    if not args.master_addr:
        args.master_addr = _first_host_address(next(iter(active)), args.ssh_port)
    if args.autotuning:
        return run_autotuning(args, active)
    active = _limit_selection(active, args.num_nodes, args.num_gpus)
    world_info_base64 = encode_world_info(active)
    # 4. command
    env = os.environ.copy()
    kill_cmd = None
    if args.force_multi or len(active) > 1:
        cmd, kill_cmd, env = _multi_node_command(args, env, resource_pool, active, world_info_base64)
    else:
        cmd = build_launch_command(args, active, world_info_base64)
    logger.info(f"cmd = {' '.join(cmd)}")
    # 5. supervise
    rc = _supervise(cmd, env, kill_cmd)
    if rc > 0:
        sys.exit(rc)


if __name__ == "__main__":
    main()
