"""Launcher logic as pure functions (the reference tests multi-node code the same way: tests/unit/launcher/*)."""
import argparse
import base64
import json

import pytest


def test_hostfile_and_filters():
    from deepspeedsyclsupport_b200.launcher import runner as R
    pool = R._parse_hostfile(["worker-0 slots=4", "# comment", "", "worker-1 slots=4"])
    assert dict(pool) == {"worker-0": 4, "worker-1": 4}
    with pytest.raises(ValueError):
        R._parse_hostfile(["worker-0 slots=4", "worker-0 slots=2"])
    with pytest.raises(ValueError):
        R._parse_hostfile(["worker-0 slotz=4"])
    act = R.parse_inclusion_exclusion(pool, "worker-1:0,2", "")
    assert dict(act) == {"worker-1": [0, 2]}
    act = R.parse_inclusion_exclusion(pool, "", "worker-0@worker-1:1")
    assert dict(act) == {"worker-1": [0, 2, 3]}
    with pytest.raises(ValueError):
        R.parse_inclusion_exclusion(pool, "worker-0", "worker-1")
    enc = R.encode_world_info({"worker-0": [0, 1]})
    assert json.loads(base64.urlsafe_b64decode(enc)) == {"worker-0": [0, 1]}
    assert R.parse_num_nodes("2:4", True) == (2, 4)


def test_rank_env_mapping():
    from deepspeedsyclsupport_b200.launcher.launch import build_rank_env
    mapping, envs = build_rank_env({"a": [0, 1], "b": [0, 1, 2]}, node_rank=1, master_addr="10.0.0.1", master_port=29500)
    assert dict(mapping) == {"a": [0, 1], "b": [2, 3, 4]}
    assert [e["RANK"] for e in envs] == ["2", "3", "4"] and envs[0]["WORLD_SIZE"] == "5"
    assert envs[2]["LOCAL_RANK"] == "2" and envs[0]["CROSS_RANK"] == "1" and envs[0]["LOCAL_SIZE"] == "3"


def _args(**kw):
    d = dict(user_script="train.py", user_args=["--x", "1"], launcher_args="", master_addr="10.0.0.1", master_port=29500, no_python=False,
             module=False, no_local_rank=False, save_pid=False, elastic_training=False, include="", exclude="", num_nodes=-1, num_gpus=-1,
             hostfile="/job/hostfile", ssh_port=None, launcher="pdsh", max_elastic_nodes=-1, min_elastic_nodes=-1, bind_cores_to_rank=False,
             bind_core_list=None, enable_each_rank_log="None")
    d.update(kw)
    return argparse.Namespace(**d)


def test_multinode_runner_commands():
    from deepspeedsyclsupport_b200.launcher import multinode_runner as M
    res = {"worker-0": 8, "worker-1": 8}
    cmd, kill, env = M.PDSHRunner(_args(), "WORLDINFO").get_cmd({}, {k: list(range(v)) for k, v in res.items()})
    assert cmd[0] == "pdsh" and "worker-0,worker-1" in cmd and "--node_rank=%n" in cmd and cmd[-3:] == ["train.py", "--x", "1"]
    assert env["PDSH_RCMD_TYPE"] == "ssh"
    cmd, _, _ = M.OpenMPIRunner(_args(launcher="openmpi"), "W", res).get_cmd({}, res)
    assert cmd[:3] == ["mpirun", "-n", "16"] and "-hostfile" in cmd
    cmd, _, _ = M.SlurmRunner(_args(launcher="slurm", num_nodes=2), "W", res).get_cmd({}, res)
    assert cmd[:3] == ["srun", "-n", "16"] and "--nodes" in cmd
    cmd, _, _ = M.IMPIRunner(_args(launcher="impi"), "W", res).get_cmd({}, res)
    assert cmd.count("-n") == 16 and "I_MPI_PIN" in cmd
    with pytest.raises(ValueError):
        M.OpenMPIRunner(_args(launcher="openmpi", include="worker-0"), "W", res)


def test_numactl_cmd():
    from deepspeedsyclsupport_b200.utils.numa import get_numactl_cmd, parse_range_list
    assert parse_range_list("0-3,8,10-11") == [0, 1, 2, 3, 8, 10, 11]
    cores, cmd = get_numactl_cmd("0-7", 2, 1)
    assert cores == 4 and cmd[0] == "numactl" and cmd[-1] == "4-7"
