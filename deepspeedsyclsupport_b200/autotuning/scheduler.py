"""Experiment execution for the autotuner.

ResourceManager keeps a queue of experiments and a pool of (host, gpu slots) nodes; an experiment reserves
`num_nodes x num_gpus` slots, runs the user command through the launcher with its own ds_config and master port, and its
result is read back from the metric file the engine writes when `autotuning.enabled` (see engine._autotuning_tick).
Experiments that fit side by side (e.g. two 4-GPU runs on one 8-GPU node) run concurrently in threads.
`run_fn` can be injected (tests, or custom cluster submitters): run_fn(exp: dict, reservations) -> None must write
exp["ds_config"]["autotuning"]["metric_path"].  API parity: deepspeed/autotuning/scheduler.py (ResourceManager, Node,
Reservation, schedule_experiments / run / parse_results)."""
import base64
import copy
import json
import os
import subprocess
import sys
import threading
import time

from ..utils.logging import logger
from .constants import AUTOTUNING, AUTOTUNING_METRIC_PATH, BUFSIZE, DS_CONFIG
from .utils import find_replace, get_val_by_key, search_error, was_interruptted

TIMEOUT = 5
thread_lock = threading.Lock()


class Node(object):

    def __init__(self, host, max_slots):
        self.host = host
        self.max_slots = max_slots
        self.idle_slots = list(range(max_slots))

    def reserve_slots(self, slot_request: int) -> list:
        if len(self.idle_slots) >= slot_request:
            return [self.idle_slots.pop(0) for _ in range(slot_request)]

    def restore_slots(self, slots: list):
        self.idle_slots += slots


class Reservation(object):

    def __init__(self, node, slots):
        self.node = node
        self.slots = slots

    def restore_slots(self):
        self.node.restore_slots(self.slots)

    def desc(self):
        return f"{self.node.host}:{','.join(str(x) for x in self.slots)}@"


class ResourceManager(object):

    def __init__(self, args, hosts, num_gpus_per_node, results_dir, exps_dir, arg_mappings, run_fn=None):
        self.results_dir = results_dir
        self.exps_dir = exps_dir
        self.nodes = [Node(host, num_gpus_per_node) for host in hosts]
        self.experiment_queue = []
        self.running_experiments = {}
        self.finished_experiments = {}
        self.experiment_count = 0
        self.exp_paths = set()
        self.args = args
        self.arg_mappings = {}
        if arg_mappings is not None:
            for k, v in arg_mappings.items():
                self.arg_mappings[k.strip()] = v.strip()
        self.run_fn = run_fn or run_experiment

    def schedule_experiments(self, exp_paths):
        for exp_path in exp_paths:
            if exp_path in self.exp_paths:
                continue
            self.exp_paths.add(exp_path)
            with open(exp_path, "r") as fd:
                exp = json.load(fd)
            exp["exp_id"] = self.experiment_count
            self.experiment_count += 1
            result_dir = exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
            if getattr(self.args, "user_script", None):
                exp["user_script"] = self.args.user_script
                exp["user_args"] = list(getattr(self.args, "user_args", []))
            exp["hostfile"] = getattr(self.args, "hostfile", None)
            exp["launcher_args"] = getattr(self.args, "launcher_args", "")
            metric_file = os.path.join(result_dir, "metrics.json")
            exp["ds_config"][AUTOTUNING][AUTOTUNING_METRIC_PATH] = metric_file
            stderr_file = os.path.join(result_dir, "stderr.log")
            model_info_file = os.path.join(result_dir, "model_info.json")
            # skip experiments that already produced a result (resume) unless they were interrupted
            if os.path.exists(result_dir) and os.path.exists(stderr_file) and not was_interruptted(stderr_file):
                err = search_error(stderr_file)
                done = os.path.exists(metric_file) or (exp["ds_config"][AUTOTUNING].get("model_info", {}) or {}).get("profile") and os.path.exists(model_info_file)
                if done or err:
                    logger.info(f"Skipping exp {exp['name']} whose result already exists")
                    self.finished_experiments[exp["exp_id"]] = (exp, err)
                    continue
            self.experiment_queue.append(exp)

    def resource_request(self, exp):
        num_nodes, num_gpus = exp["num_nodes"], exp["num_gpus"]
        reservations = []
        for node in self.nodes:
            if num_nodes == 0:
                break
            slots = node.reserve_slots(slot_request=num_gpus)
            if slots:
                reservations.append(Reservation(node=node, slots=slots))
                num_nodes -= 1
        if num_nodes == 0:
            return reservations
        for r in reservations:
            r.restore_slots()
        return None

    def status(self):
        return ", ".join(f"{node.host} ({len(node.idle_slots)} idle gpus)" for node in self.nodes)

    def experiment_check(self, pbar=None):
        finished = []
        for exp_id, (thread, exp_json, reservations, start_time) in self.running_experiments.items():
            thread.join(timeout=TIMEOUT)
            if not thread.is_alive():
                err = search_error(os.path.join(exp_json["result_dir"], "stderr.log")) if not os.path.exists(
                    exp_json["ds_config"][AUTOTUNING][AUTOTUNING_METRIC_PATH]) else None
                finished.append((exp_id, err))
                self.finished_experiments[exp_id] = (exp_json, err)
                logger.info(f"Finished running exp_id = {exp_id} in {time.time() - start_time:.1f}s")
                for r in reservations:
                    r.restore_slots()
        for exp_id, _ in finished:
            del self.running_experiments[exp_id]

    def run_job(self, exp: dict, reservations):
        exp["master_port"] = int(getattr(self.args, "master_port", 29500)) + exp["exp_id"]
        exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
        user_args = exp.get("user_args", [])
        # values of tuned ds_config keys that the script takes on its command line (--per_device_train_batch_size etc.)
        for key, val in self.arg_mappings.items():
            nval = get_val_by_key(exp, key)
            if nval is not None and str(nval) != "auto":
                if val in user_args:
                    user_args[user_args.index(val) + 1] = str(nval)
                else:
                    user_args += [val, str(nval)]
        exp["user_args"] = user_args
        t = threading.Thread(target=self.run_fn, args=(exp, reservations, getattr(self.args, "user_script", None), user_args))
        t.start()
        self.running_experiments[exp["exp_id"]] = (t, exp, reservations, time.time())

    def run(self):
        while len(self.experiment_queue) > 0:
            exp = self.experiment_queue.pop(0)
            logger.info(f'Checking resource availability for {exp["name"]}: {self.status()}')
            reservations = self.resource_request(exp)
            if not reservations:
                self.experiment_queue.insert(0, exp)
                if not self.running_experiments:
                    raise RuntimeError(f"experiment {exp['name']} asks for more resources than the pool has")
                self.experiment_check()
            else:
                self.run_job(exp, reservations)
        while len(self.running_experiments) > 0:
            self.experiment_check()

    def save_exp_results_to_database(self, message, ranks=None, path=None):
        if path is None:
            return
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "a") as fd:
            fd.write(json.dumps(message) + "\n")

    def parse_results(self, metric):
        """-> (best experiment dict, best metric value) over the finished experiments ('latency' is minimised)."""
        max_throughput = sys.float_info.min
        best_exp_id = -1
        for exp_id, (exp, err) in self.finished_experiments.items():
            if err:
                logger.info(f"The experiment exp_id = {exp_id}, exp_name = {exp['name']}, did not run successfully with error = {err}")
                continue
            metric_file = exp[DS_CONFIG][AUTOTUNING][AUTOTUNING_METRIC_PATH]
            if os.path.exists(metric_file):
                with open(metric_file, "r") as f:
                    results = json.load(f)
                    curr = results[metric] if metric != "latency" else -results[metric]
                    if curr > max_throughput:
                        max_throughput = curr
                        best_exp_id = exp_id
                    exp["results"] = results
        if best_exp_id != -1:
            best_exp, _ = self.finished_experiments[best_exp_id]
            return best_exp, max_throughput if metric != "latency" else -max_throughput
        return None, None

    def clear(self):
        self.experiment_queue = []
        for exp_id, (thread, exp_json, reservations, start_time) in self.running_experiments.items():
            thread.join(timeout=TIMEOUT)
            clean_up(exp_json, reservations)
        self.running_experiments = {}
        self.finished_experiments = {}
        self.exp_paths = set()


def run_experiment(exp: dict, reservations, user_script, user_args):
    """Default runner: `python -m <launcher> --include host:slots ... user_script args` with the experiment's ds_config."""
    include_str = "".join(r.desc() for r in reservations)[:-1]
    master_port = exp["master_port"]
    exp_dir = exp["result_dir"]
    os.makedirs(exp_dir, exist_ok=True)
    exp["ds_config_path"] = os.path.join(exp_dir, "ds_config.json")
    exp["ds_config_base64"] = base64.urlsafe_b64encode(json.dumps(exp["ds_config"]).encode("utf-8")).decode("utf-8")
    user_args = list(user_args)
    # point --deepspeed_config / --deepspeed at this experiment's config
    for flag in ("--deepspeed_config", "--deepspeed"):
        if flag in user_args:
            idx = user_args.index(flag)
            if idx + 1 < len(user_args) and not user_args[idx + 1].startswith("--"):
                user_args[idx + 1] = exp["ds_config_path"]
    with open(exp["ds_config_path"], "w", buffering=BUFSIZE) as fd:
        json.dump(exp["ds_config"], fd)
        fd.flush()
        os.fsync(fd)
    with open(os.path.join(exp_dir, "exp.json"), "w", buffering=BUFSIZE) as fd:
        json.dump(exp, fd)
        fd.flush()
        os.fsync(fd)
    cmd = [sys.executable, "-m", "deepspeedsyclsupport_b200.launcher.runner"]
    if exp.get("hostfile") and os.path.isfile(str(exp["hostfile"])):
        cmd += ["--hostfile", exp["hostfile"]]
    cmd += ["--include", include_str, "--master_port", str(master_port)]
    if exp.get("launcher_args"):
        cmd += ["--launcher_args", exp["launcher_args"]]
    cmd += [user_script] + user_args
    with open(os.path.join(exp_dir, "cmd.txt"), "w", buffering=BUFSIZE) as fd:
        fd.write(" ".join(cmd) + "\n")
    logger.info(f"Launching exp_id = {exp['exp_id']}, exp_name = {exp['name']}, with resource = {include_str}, and ds_config = {os.path.abspath(exp['ds_config_path'])}")
    with open(os.path.join(exp_dir, "stdout.log"), "wb") as out, open(os.path.join(exp_dir, "stderr.log"), "wb") as err:
        result = subprocess.Popen(cmd, stdout=out, stderr=err)
        result.wait()
        out.flush()
        err.flush()
    clean_up(exp, reservations)
    logger.info(f"Done running exp_id = {exp['exp_id']}, exp_name = {exp['name']}, with resource = {include_str}")


def clean_up(exp: dict, reservations):
    """Nothing may be left running on the reserved slots.  The launcher tears down its own process tree when it exits
    (launch.py kills its children on failure/termination), so there is no stray-process sweep here."""
    return
