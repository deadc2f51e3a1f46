"""Multi-node command builders.  Each runner turns (args, world info, exported env) into the argv of ONE command that,
run on the launch host, starts the job on every node:

    pdsh     one `launcher.launch` per node over ssh; launch.py then forks the local ranks
    openmpi / mpich / impi / mvapich     one process per rank directly under mpirun (ranks read their ids from MPI env)
    slurm    srun -n <ranks>

`get_cmd(environment, active_resources)` returns (argv, kill_argv, environment).  Everything here is pure string
assembly so it is unit-tested without any of the launchers installed.
(CLI/flag parity with deepspeed/launcher/multinode_runner.py.)"""
import os
import shutil
import subprocess
import sys
import warnings
from abc import ABC, abstractmethod
from shlex import split

from ..utils.logging import logger
from .constants import MVAPICH_TMP_HOSTFILE, PDSH_MAX_FAN_OUT

LAUNCH_MODULE = "deepspeedsyclsupport_b200.launcher.launch"


class MultiNodeRunner(ABC):
    name = "base"
    # runners that start ranks directly under MPI cannot honour per-node include/exclude or --num_nodes/--num_gpus
    supports_resource_filters = True

    def __init__(self, args, world_info_base64, resource_pool=None):
        self.args = args
        self.resource_pool = resource_pool
        self.validate_args()
        self.user_arguments = self.parse_user_args()
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.exports = {}

    @abstractmethod
    def backend_exists(self):
        """Is the launcher binary available on this host?"""

    @abstractmethod
    def get_cmd(self, environment, active_resources):
        """(argv, kill_argv, environment)"""

    def add_export(self, key, var):
        self.exports[key.strip()] = var.strip()

    def parse_user_args(self):
        return self.args.user_args

    def validate_args(self):
        if self.supports_resource_filters:
            return
        if self.args.include != "" or self.args.exclude != "":
            raise ValueError(f"{self.name} backend does not support worker include/exclusion")
        if self.args.num_nodes != -1 or self.args.num_gpus != -1:
            raise ValueError(f"{self.name} backend does not support limiting num nodes/gpus")

    # ---- shared fragments
    def _python(self):
        """Interpreter prefix for the user script: [] with --no_python, else python -u [-m]."""
        if self.args.no_python:
            return []
        return [sys.executable, "-u"] + (["-m"] if self.args.module else [])

    def _user_command(self):
        return [self.user_script] + list(self.user_arguments)

    def _uniform_ranks_per_node(self, what):
        per_node = list(self.resource_pool.values())
        if any(n != per_node[0] for n in per_node):
            raise ValueError(f"{what} requires same number of devices per node")
        return per_node[0], sum(per_node)

    def _rendezvous(self, flag, total, per_node):
        pairs = (("MASTER_ADDR", self.args.master_addr), ("MASTER_PORT", self.args.master_port), ("WORLD_SIZE", total), ("LOCAL_SIZE", per_node))
        out = []
        for k, v in pairs:
            out += [flag, k, str(v)]
        return out


class PDSHRunner(MultiNodeRunner):
    name = "pdsh"

    def backend_exists(self):
        return shutil.which("pdsh")

    def parse_user_args(self):
        # JSON blobs must survive the remote shell
        return [f"'{a}'" if (a.startswith("{") and a.endswith("}")) else a for a in self.args.user_args]

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        if self.args.ssh_port is not None:
            environment["PDSH_SSH_ARGS_APPEND"] = f"{environment.get('PDSH_SSH_ARGS_APPEND', '')} -p {self.args.ssh_port}"
        workers = ",".join(active_resources.keys())
        logger.info("Running on the following workers: %s" % workers)
        pdsh = ["pdsh", "-S", "-f", str(PDSH_MAX_FAN_OUT), "-w", workers] + split(self.args.launcher_args)
        exports = "".join(f"export {k}={v}; " for k, v in self.exports.items())
        # %n is replaced by pdsh with the node's index in the -w list = its node rank
        launch = [exports, f"cd {os.path.abspath('.')};", sys.executable, "-u", "-m", LAUNCH_MODULE, f"--world_info={self.world_info_base64}",
                  "--node_rank=%n", f"--master_addr={self.args.master_addr}", f"--master_port={self.args.master_port}"]
        for flag, on in (("--no_python", self.args.no_python), ("--module", self.args.module), ("--no_local_rank", self.args.no_local_rank)):
            if on:
                launch.append(flag)
        # every node launcher records its PID under /tmp/<runner pid>.deepspeed so the kill command can signal exactly
        # that process (which tears down its own children) instead of matching command lines by pattern
        launch += ["--save_pid", f"{os.getpid()}"]
        if self.args.elastic_training:
            launch += ["--enable_elastic_training", f"--max_elastic_nodes={self.args.max_elastic_nodes}", f"--min_elastic_nodes={self.args.min_elastic_nodes}"]
        pid_file = f"/tmp/{os.getpid()}.deepspeed"
        kill = pdsh + [f"test -f {pid_file} && kill -TERM $(cat {pid_file})"]
        return pdsh + launch + self._user_command(), kill, environment


class OpenMPIRunner(MultiNodeRunner):
    name = "openmpi"
    supports_resource_filters = False

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)
        self.add_export("UCX_TLS", "tcp")

    def backend_exists(self):
        return shutil.which("ompi_info")

    def get_cmd(self, environment, active_resources):
        total = sum(self.resource_pool.values())
        mpirun = ["mpirun", "-n", f"{total}", "-hostfile", f"{self.args.hostfile}", "--mca", "btl", "^openib", "--mca", "btl_tcp_if_include", "eth0"]
        mpirun += split(self.args.launcher_args)
        for k, v in self.exports.items():
            mpirun += ["-x", f"{k}={v}"]
        return mpirun + self._python() + self._user_command(), [], environment


class MPICHRunner(MultiNodeRunner):
    name = "mpich"
    supports_resource_filters = False

    def backend_exists(self):
        return shutil.which("mpirun")

    def get_cmd(self, environment, active_resources):
        per_node, total = self._uniform_ranks_per_node("MPICH")
        mpirun = ["mpirun"] + split(self.args.launcher_args)
        for k, v in self.exports.items():
            mpirun += ["-genv", f"{k}", f"{v}"]
        mpirun += self._rendezvous("-genv", total, per_node)
        mpirun += ["-hosts", ",".join(self.resource_pool.keys())]
        # MPICH exports PMI_RANK etc.; the helper maps them to RANK / LOCAL_RANK before running the script
        helper = [os.path.join(os.path.dirname(os.path.realpath(__file__)), "launcher_helper.py"), "--launcher", self.args.launcher]
        if self.args.no_python:
            helper.append("--no_python")
        elif self.args.module:
            helper.append("--module")
        return mpirun + self._python() + helper + self._user_command(), [], environment


class IMPIRunner(MultiNodeRunner):
    name = "impi"
    supports_resource_filters = False

    def backend_exists(self):
        return shutil.which("mpirun")

    def get_cmd(self, environment, active_resources):
        per_node, total = self._uniform_ranks_per_node("Intel MPI")
        mpirun = ["mpirun", "-ppn", f"{per_node}"] + split(self.args.launcher_args)
        for k, v in self.exports.items():
            mpirun += ["-genv", f"{k}", f"{v}"]
        if self.args.bind_cores_to_rank:
            from ..utils.numa import get_numactl_cmd
            cores_per_rank, _ = get_numactl_cmd(self.args.bind_core_list, per_node, 0)
            mpirun += ["-genv", "OMP_NUM_THREADS", str(cores_per_rank)]
        mpirun += self._rendezvous("-genv", total, per_node)
        mpirun += ["-genv", "I_MPI_PIN", "0", "-hosts", ",".join(self.resource_pool.keys())]
        # MPMD form: one `-n 1 -env RANK r -env LOCAL_RANK l <cmd>` block per rank, joined by ':'
        blocks = []
        for rank in range(total):
            blocks.append(["-n", "1", "-env", "RANK", str(rank), "-env", "LOCAL_RANK", str(rank % per_node)] + self._python() + self._user_command())
        argv = list(mpirun)
        for i, b in enumerate(blocks):
            argv += ([":"] if i else []) + b
        return argv, [], environment


class SlurmRunner(MultiNodeRunner):
    name = "slurm"

    def backend_exists(self):
        return shutil.which("sinfo")

    def get_cmd(self, environment, active_resources):
        assert not getattr(self.args, "detect_nvlink_pairs", False), "slurm backend does not support remapping visible devices"
        srun = ["srun", "-n", f"{sum(self.resource_pool.values())}"] + split(self.args.launcher_args)
        if getattr(self.args, "slurm_comment", ""):
            srun += ["--comment", self.args.slurm_comment]
        for flag, value, on in (("--include", self.args.include, self.args.include != ""), ("--exclude", self.args.exclude, self.args.exclude != ""),
                                ("--nodes", self.args.num_nodes, self.args.num_nodes > 0), ("--gpus", self.args.num_gpus, self.args.num_gpus > 0)):
            if on:
                srun += [flag, f"{value}"]
        exports = "--export=ALL" + "".join(f",{k}={v}" for k, v in self.exports.items())
        return srun + [exports, sys.executable, "-u"] + self._user_command(), [], environment


class MVAPICHRunner(MultiNodeRunner):
    name = "mvapich"
    supports_resource_filters = False
    _DEFAULT_EXPORTS = {
        "MV2_SMP_USE_CMA": "0",  # CMA is unavailable in most containers
        "MV2_DEBUG_SHOW_BACKTRACE": "1",
        "MV2_USE_CUDA": "1",  # CUDA-aware MPI
        "MV2_SUPPORT_DL": "1",
        "MV2_ENABLE_AFFINITY": "0",
        "MV2_INTER_ALLGATHER_TUNING": "5",
        "MV2_CUDA_USE_NAIVE": "0",
    }

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64, resource_pool)
        for k, v in self._DEFAULT_EXPORTS.items():
            self.add_export(k, v)

    def backend_exists(self):
        if not shutil.which("mpiname"):
            warnings.warn("mpiname does not exist, mvapich is not installed properly")
            return False
        found = subprocess.check_output("mpiname", shell=True).decode("utf-8").strip()
        if "MVAPICH2-GDR" in found:
            return True
        warnings.warn(f"Expected MVAPICH2-GDR as return for mpiname but received {found}")
        return False

    def get_cmd(self, environment, active_resources):
        per_node, total = self._uniform_ranks_per_node("mvapich")
        with open(MVAPICH_TMP_HOSTFILE, "w") as fd:
            fd.writelines(f"{host}\n" for host in self.resource_pool.keys())
        mpirun = ["mpirun", "-np", f"{total}", "-ppn", f"{per_node}", "--hostfile", f"{MVAPICH_TMP_HOSTFILE}"] + split(self.args.launcher_args)
        for k, v in self.exports.items():
            mpirun += ["-env", f"{k}={v}"]
        return mpirun + self._python() + self._user_command(), [], environment
