"""NUMA core binding helpers for the launcher (parity: deepspeed/utils/numa.py)."""
import os
import subprocess


def get_numa_cores():
    ret = []
    try:
        output = subprocess.check_output(["numactl", "--hardware"]).decode("utf-8")
    except Exception:
        return [list(range(os.cpu_count() or 1))]
    for line in output.split("\n"):
        line = line.strip()
        if line.startswith("available:"):
            continue
        if "cpus:" in line and line.startswith("node"):
            ret.append([int(c) for c in line.split("cpus:")[1].split()])
    return ret or [list(range(os.cpu_count() or 1))]


def parse_range(rng):
    try:
        value = int(rng)
        return range(value, value + 1)
    except ValueError:
        parts = rng.split("-")
        if len(parts) != 2:
            raise ValueError("Bad range: '%s', range must be either a number or two number separated by dash" % (rng, ))
        start, end = int(parts[0]), int(parts[1])
        if start > end:
            raise ValueError("Bad range: '%s', range end must larger than or equal to start" % (rng, ))
        return range(start, end + 1)


def parse_range_list(range_str):
    number_list = []
    last = -1
    for sub in range_str.split(","):
        sub_number_list = parse_range(sub)
        if sub_number_list[0] <= last:
            raise ValueError("Bad range: '%s', sub ranges must not overlap with each other and should be in ascend order" % (range_str, ))
        last = sub_number_list[-1]
        number_list.extend(sub_number_list)
    return number_list


def get_numactl_cmd(bind_core_list, num_local_procs, local_rank):
    numactl_cmd = []
    if bind_core_list is not None:
        core_list = parse_range_list(bind_core_list)
        total_cores = len(core_list)
    else:
        total_cores = os.cpu_count() or 1
        core_list = list(range(total_cores))
    cores_per_rank = max(1, total_cores // num_local_procs)
    core_list_for_rank = core_list[cores_per_rank * local_rank:cores_per_rank * (local_rank + 1)]
    numactl_cmd.append("numactl")
    numa_cores = get_numa_cores()
    for i, cores in enumerate(numa_cores):
        if set(core_list_for_rank) <= set(cores):
            numactl_cmd.append("-m")
            numactl_cmd.append(f"{i}")
            break
    core_list_str = f"{core_list_for_rank[0]}"
    last_core = core_list_for_rank[0]
    first_core = last_core
    for core_id in core_list_for_rank[1:]:
        if core_id == last_core + 1:
            last_core = core_id
            continue
        else:
            if first_core != last_core:
                core_list_str = f"{core_list_str}-{last_core}"
            core_list_str = f"{core_list_str},{core_id}"
            first_core = last_core = core_id
    if first_core != last_core:
        core_list_str = f"{core_list_str}-{last_core}"
    numactl_cmd.append("-C")
    numactl_cmd.append(f"{core_list_str}")
    return cores_per_rank, numactl_cmd
