"""Helper executed by mpirun-style launchers: maps the MPI rank variables onto RANK / LOCAL_RANK and runs the script
(parity: deepspeed/launcher/launcher_helper.py)."""
import argparse
import os
import subprocess
import sys


def parse_args(args=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--launcher", default="mpich", type=str)
    parser.add_argument("--module", action="store_true")
    parser.add_argument("--no_python", action="store_true")
    parser.add_argument("user_script", type=str)
    parser.add_argument("user_args", nargs=argparse.REMAINDER)
    return parser.parse_args(args=args)


def env_mapping(env, rank_name_list=None, local_rank_name_list=None):
    rank = None
    for rank_name in rank_name_list:
        if rank_name in env:
            if rank is None:
                rank = env.get(rank_name)
            elif rank != env.get(rank_name):
                raise EnvironmentError("rank number doesn't match!")
    if rank is None:
        raise EnvironmentError("rank number is not in current env!")
    env["RANK"] = rank
    local_rank = None
    for local_rank_name in local_rank_name_list:
        if local_rank_name in env:
            if local_rank is None:
                local_rank = env.get(local_rank_name)
            elif local_rank != env.get(local_rank_name):
                raise EnvironmentError("local_rank number doesn't match!")
    if local_rank is None:
        raise EnvironmentError("rank number is not in current env!")
    env["LOCAL_RANK"] = local_rank
    return env


def main(args=None):
    args = parse_args(args)
    env = os.environ.copy()
    args.launcher = args.launcher.lower()
    if args.launcher == "mpich":
        rank_name_list, local_rank_name_list = ["PMIX_RANK"] + ["PMI_RANK"], ["PALS_LOCAL_RANKID", "MPI_LOCALRANKID"]
        env = env_mapping(env, rank_name_list=rank_name_list, local_rank_name_list=local_rank_name_list)
    else:
        raise NotImplementedError(f"Unknown launcher {args.launcher}")
    python_exec = []
    if not args.no_python:
        python_exec += [sys.executable, "-u"]
        if args.module:
            python_exec.append("-m")
    cmd = python_exec + [args.user_script] + args.user_args
    result = subprocess.Popen(cmd, env=env)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
