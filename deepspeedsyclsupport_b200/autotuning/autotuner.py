"""Autotuner: finds the ZeRO stage / micro-batch size / ZeRO knobs that maximise throughput (or minimise latency) for
the user's training script on the available GPUs, by running short profiling experiments.

Flow of tune():
  1. model-info run   -- one experiment with the minimum-memory config (ZeRO-3, mbs 1) and `model_info.profile`: the
                         engine reports parameter count and activation memory per GPU, then exits.
  2. per ZeRO stage (those the user allowed and whose estimated resident memory fits the GPU):
       a. micro-batch sweep: mbs = min, 2*min, 4*min ... until an experiment fails (OOM) or the cap is reached; the
          sweep gives the feasible range and its best point;
       b. `num_tuning_micro_batch_sizes` candidates around the best point are combined with the stage's tuning space
          (bucket sizes, overlap_comm ...) unless `fast` mode is on, and searched by the configured tuner
          (gridsearch / random / model_based) with early stopping.
     A later stage is only explored if it can reach a larger micro-batch or a better metric than the previous one.
  3. the best experiment overall is written to <results_dir>/ds_config_optimal.json + cmd_optimal.txt and summarised.
Memory model (bytes per GPU, P parameters, N data-parallel ranks, mixed precision): params 2P, grads 2P, optimizer 12P
(fp32 master + two moments) -- stage 1 shards the optimizer, stage 2 also the gradients, stage 3 everything; activation
memory scales linearly with the micro-batch size from the measured mbs=1 value.
CLI / file-layout parity: deepspeed/autotuning/autotuner.py (results_dir, exps_dir, profile_model_info, summary table,
ds_config_optimal.json, cmd_optimal.txt, --autotuning run)."""
import copy
import datetime
import json
import os
import shutil
import subprocess
import sys
import time

from ..runtime.config_utils import dict_raise_error_on_duplicate_keys
from ..runtime.constants import GRADIENT_ACCUMULATION_STEPS, TRAIN_BATCH_SIZE, TRAIN_MICRO_BATCH_SIZE_PER_GPU
from ..utils.logging import logger
from .config import DeepSpeedAutotuningConfig
from .constants import (AUTOTUNING, AUTOTUNING_METRIC_LATENCY, AUTOTUNING_METRIC_PATH, AUTOTUNING_MODEL_INFO_PATH, AUTOTUNING_TUNER_GRIDSEARCH,
                        AUTOTUNING_TUNER_MODELBASED, AUTOTUNING_TUNER_RANDOM, DEFAULT_MIN_MEM_CONFIG, DEFAULT_TUNING_SPACES, DS_CONFIG,
                        METRIC_PERCENT_DIFF_CONST, MODEL_INFO, MODEL_INFO_NUM_PARAMS, MODEL_INFO_PROFILE)
from .scheduler import ResourceManager
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner
from .utils import (canonical_name, get_all_configs, get_first_config, get_tuning_keys, memory_to_string, number_to_string, prune_configs,
                    replace_dict, write_experiments)

try:
    from tabulate import tabulate
except ImportError:
    tabulate = None

ZERO_OPTIMIZATION = "zero_optimization"
ZERO_OPTIMIZATION_STAGE = "stage"


class Autotuner(object):

    def __init__(self, args, active_resources, run_fn=None):
        self.args = args
        self.selected_exp_dir = None
        self.user_config = self._get_user_config(args.user_args)
        assert self.user_config is not None, "DeepSpeed configuration is not provided"
        self.autotuning_config = DeepSpeedAutotuningConfig(**(self.user_config.get(AUTOTUNING, {}) or {}))
        if self.user_config.get(AUTOTUNING):
            if AUTOTUNING_METRIC_PATH in self.user_config[AUTOTUNING]:
                del self.user_config[AUTOTUNING][AUTOTUNING_METRIC_PATH]
        self.results_dir = self.autotuning_config.results_dir
        self.exps_dir = self.autotuning_config.exps_dir
        for d in (self.exps_dir, self.results_dir):
            if self.autotuning_config.overwrite and os.path.exists(d):
                shutil.rmtree(d, ignore_errors=True)
            os.makedirs(d, exist_ok=True)
        self.exp_num_nodes = len(active_resources)
        self.exp_num_gpus = min(len(v) for v in active_resources.values()) if active_resources else 0
        if getattr(args, "num_nodes", -1) and args.num_nodes > 0:
            self.exp_num_nodes = args.num_nodes
        if getattr(args, "num_gpus", -1) and args.num_gpus > 0:
            self.exp_num_gpus = args.num_gpus
        assert self.exp_num_nodes <= len(active_resources) or not active_resources, "more nodes requested than available"
        self.rm = ResourceManager(args=args, hosts=list(active_resources.keys()), num_gpus_per_node=self.exp_num_gpus, results_dir=self.results_dir,
                                  exps_dir=self.exps_dir, arg_mappings=self.autotuning_config.arg_mappings, run_fn=run_fn)
        self.records = {}
        self.optimal_cmd = None
        self.optimal_ds_config = None
        self.model_info = None
        self.start_time = None

    # ------------------------------------------------------------------ configuration
    def _get_user_config(self, user_args):
        """The ds_config the script is launched with: the path following --deepspeed_config or --deepspeed."""
        path = None
        for flag in ("--deepspeed_config", "--deepspeed"):
            if flag in user_args:
                idx = user_args.index(flag)
                if idx + 1 < len(user_args) and not user_args[idx + 1].startswith("--"):
                    path = user_args[idx + 1]
                    break
        if path is None:
            return None
        assert os.path.isfile(path), f"DeepSpeed configuration file: {path} is not an existing file"
        logger.debug(f"user_config_file = {path}")
        with open(path, "r") as f:
            return json.load(f, object_pairs_hook=dict_raise_error_on_duplicate_keys)

    def metric(self):
        return self.autotuning_config.metric

    def fast_enabled(self):
        return self.autotuning_config.fast

    def max_train_batch_size(self):
        return self.autotuning_config.max_train_batch_size

    def mp_size(self):
        return self.autotuning_config.mp_size

    def max_train_micro_batch_size_per_gpu(self):
        cap = self.autotuning_config.max_train_micro_batch_size_per_gpu
        if self.max_train_batch_size() and self.max_train_batch_size() > 0:
            cap = min(cap, self.max_train_batch_size() * self.mp_size() // max(1, self.exp_num_gpus * self.exp_num_nodes))
        return max(cap, 1)

    def min_train_micro_batch_size_per_gpu(self):
        return self.autotuning_config.min_train_micro_batch_size_per_gpu

    def num_tuning_micro_batch_sizes(self):
        return self.autotuning_config.num_tuning_micro_batch_sizes

    def fp16_enabled(self):
        return bool(self.user_config.get("fp16", {}).get("enabled", False) or self.user_config.get("bf16", {}).get("enabled", False))

    def get_gas_from_user_config(self):
        gas = self.user_config.get(GRADIENT_ACCUMULATION_STEPS, 1)
        return 1 if gas == "auto" else int(gas)

    # ------------------------------------------------------------------ memory model
    def get_gpu_memory_info(self):
        try:
            import torch
            if torch.cuda.is_available():
                return torch.cuda.get_device_properties(0).total_memory
        except Exception:
            pass
        return 180 * 1024**3  # B200

    def get_activation_memory_per_gpu(self):
        return int((self.model_info or {}).get("activation_mem_per_gpu", 0) or 0)

    def get_model_num_params(self):
        return int((self.model_info or {}).get(MODEL_INFO_NUM_PARAMS, 0) or 0)

    def get_instantiation_memory_required_per_gpu(self, zero_stage):
        P = self.get_model_num_params()
        n = max(1, self.exp_num_gpus * self.exp_num_nodes // self.mp_size())
        half = self.fp16_enabled()
        params = P * (2 if half else 4)
        grads = P * (2 if half else 4)
        optim = P * (12 if half else 8)  # fp32 master + exp_avg + exp_avg_sq (no separate master in pure fp32)
        if zero_stage >= 1:
            optim /= n
        if zero_stage >= 2:
            grads /= n
        if zero_stage >= 3:
            params /= n
        return (params + grads + optim) / self.mp_size()

    # ------------------------------------------------------------------ experiments
    def _exp(self, name, ds_config):
        cfg = copy.deepcopy(ds_config)
        at = cfg.setdefault(AUTOTUNING, {})
        at["enabled"] = True
        at.setdefault("start_profile_step", self.autotuning_config.start_profile_step)
        at.setdefault("end_profile_step", self.autotuning_config.end_profile_step)
        return {"name": name, DS_CONFIG: cfg, "num_gpus": self.exp_num_gpus, "num_nodes": self.exp_num_nodes}

    def _base_config(self, space, mbs):
        """user config overlaid with the first point of `space`, at micro-batch `mbs` (gas kept, batch recomputed)."""
        cfg = copy.deepcopy(self.user_config)
        replace_dict(cfg, get_first_config(space) or {}, ignored_keys=[])
        gas = self.get_gas_from_user_config()
        cfg[TRAIN_MICRO_BATCH_SIZE_PER_GPU] = mbs
        cfg[GRADIENT_ACCUMULATION_STEPS] = gas
        cfg[TRAIN_BATCH_SIZE] = mbs * gas * self.exp_num_gpus * self.exp_num_nodes // self.mp_size()
        return cfg

    def _run(self, exps):
        """Run a list of experiment dicts; returns {name: (exp, metric value or None)}."""
        paths = write_experiments(exps, self.exps_dir)
        self.rm.schedule_experiments(paths)
        self.rm.run()
        out = {}
        for exp_id, (exp, err) in self.rm.finished_experiments.items():
            val = None
            mfile = exp[DS_CONFIG][AUTOTUNING].get(AUTOTUNING_METRIC_PATH)
            if not err and mfile and os.path.exists(mfile):
                with open(mfile) as f:
                    results = json.load(f)
                exp["results"] = results
                val = results.get(self.metric())
            out[exp["name"]] = (exp, val)
        self.rm.clear()
        return out

    def run_ds_config(self, ds_config, exp_name):
        res = self._run([self._exp(exp_name, ds_config)])
        exp, val = res.get(exp_name, (None, None))
        return exp, val

    def model_info_profile_run(self):
        """Parameter count + activation memory of the model, from a profile run (or from the user's `model_info`)."""
        given = self.autotuning_config.model_info
        if given and MODEL_INFO_NUM_PARAMS in given and not given.get(MODEL_INFO_PROFILE, False):
            return given
        logger.info("Starting model info profile run.")
        cfg = copy.deepcopy(self.user_config)
        replace_dict(cfg, DEFAULT_MIN_MEM_CONFIG)
        cfg[TRAIN_BATCH_SIZE] = cfg[TRAIN_MICRO_BATCH_SIZE_PER_GPU] * self.exp_num_gpus * self.exp_num_nodes // self.mp_size()
        cfg[GRADIENT_ACCUMULATION_STEPS] = 1
        model_info_path = os.path.join(self.results_dir, "profile_model_info", "model_info.json")
        cfg.setdefault(AUTOTUNING, {})
        cfg[AUTOTUNING].update({"enabled": True, AUTOTUNING_MODEL_INFO_PATH: model_info_path, MODEL_INFO: {MODEL_INFO_PROFILE: True}})
        exp = self._exp("profile_model_info", cfg)
        paths = write_experiments([exp], self.exps_dir)
        self.rm.schedule_experiments(paths)
        self.rm.run()
        for exp_id, (e, err) in self.rm.finished_experiments.items():
            if err and not os.path.exists(model_info_path):
                logger.error(f"The model is not runnable with DeepSpeed with error = {err}")
                self.rm.clear()
                return None
        self.rm.clear()
        if os.path.exists(model_info_path):
            with open(model_info_path, "r") as f:
                return json.load(f)
        return None

    def _better(self, new, old):
        if new is None:
            return False
        if old is None:
            return True
        if self.metric() == AUTOTUNING_METRIC_LATENCY:
            return new < old * (1 - METRIC_PERCENT_DIFF_CONST)
        return new > old * (1 + METRIC_PERCENT_DIFF_CONST)

    def update_records(self, space_name, exp, metric_val, num_exps):
        self.records.setdefault(space_name, []).append((exp, metric_val, num_exps))

    def get_best_space_record(self, space_name):
        best, total = None, 0
        for exp, val, n in self.records.get(space_name, []):
            total += n
            if val is not None and (best is None or (val < best[1] if self.metric() == AUTOTUNING_METRIC_LATENCY else val > best[1])):
                best = (exp, val)
        return (best[0], best[1], total) if best else None

    def get_best_space_records(self):
        out = {}
        overall = None
        for name in self.records:
            rec = self.get_best_space_record(name)
            if rec:
                out[name] = rec
                if overall is None or (rec[1] < overall[1] if self.metric() == AUTOTUNING_METRIC_LATENCY else rec[1] > overall[1]):
                    overall = rec
        if overall:
            out["global"] = (overall[0], overall[1], sum(r[2] for k, r in out.items() if k != "global"))
        return out

    # ------------------------------------------------------------------ micro-batch sweep + space search
    def get_tuning_micro_batch_size_list(self, min_micro_batch_size, max_micro_batch_size, num_tuning_micro_batch_sizes):
        """Evenly spread candidates in [min, max] (always containing both ends), deduplicated, ascending."""
        if min_micro_batch_size <= 0 or max_micro_batch_size <= 0 or max_micro_batch_size < min_micro_batch_size:
            return []
        n = max(1, num_tuning_micro_batch_sizes)
        if n == 1 or min_micro_batch_size == max_micro_batch_size:
            return [max_micro_batch_size]
        step = (max_micro_batch_size - min_micro_batch_size) / (n - 1)
        return sorted({max(1, int(round(min_micro_batch_size + i * step))) for i in range(n)})

    def get_plateau_mbs(self, tuning_space_name):
        """Largest micro-batch size before the metric stops improving (< 5 % gain or a drop) in a space's records."""
        recs = [(e[DS_CONFIG][TRAIN_MICRO_BATCH_SIZE_PER_GPU], v) for e, v, _ in self.records.get(tuning_space_name, []) if v is not None]
        prev_val, plateau = None, 0
        for mbs, val in sorted(recs):
            if prev_val is not None and (val < prev_val or (val - prev_val) / prev_val < 0.05):
                break
            prev_val, plateau = val, mbs
        return plateau

    def get_val_from_user_args(self, ds_name):
        """Value the user script received on its command line for the argument mapped to `ds_name` (autotuning.arg_mappings)."""
        mappings = self.autotuning_config.arg_mappings
        user_args = list(getattr(self.args, "user_args", []) or [])
        if mappings and ds_name in mappings and mappings[ds_name] in user_args:
            i = user_args.index(mappings[ds_name])
            if i + 1 < len(user_args) and str(user_args[i + 1]).isnumeric():
                return user_args[i + 1]
        return None

    def get_min_max_micro_batch_size(self, stage, min_micro_batch_size, calculated_max_micro_batch_size):
        """(smallest runnable, largest runnable) micro-batch size of a ZeRO stage: probe the minimum, then double up to
        the calculated maximum and bisect between the last success and the first failure."""
        space, name = self._user_overlaid_space(DEFAULT_TUNING_SPACES[stage]), f"z{stage}"

        def runs(mbs):
            cfg = self._base_config(space, mbs)
            exp, val = self.run_ds_config(cfg, canonical_name(cfg, tuning_keys=[], prefix=f"{name}_probe"))
            self.update_records(name, exp, val, 1)
            return val is not None

        if not runs(min_micro_batch_size):
            return -1, -1
        lo, hi = min_micro_batch_size, None
        mbs = min_micro_batch_size * 2
        while mbs <= calculated_max_micro_batch_size:
            if runs(mbs):
                lo = mbs
                mbs *= 2
            else:
                hi = mbs
                break
        while hi is not None and hi - lo > 1:
            mid = (lo + hi) // 2
            if runs(mid):
                lo = mid
            else:
                hi = mid
        return min_micro_batch_size, lo

    def run_tuning_micro_batch_sizes(self, tuning_micro_batch_sizes, max_train_batch_size_per_gpu, tuning_micro_batch_sizes_overwritten,
                                     stage, tuning_space_name):
        """Run one experiment per candidate micro-batch size (gradient accumulation chosen to respect the batch-size cap);
        returns the largest size that ran."""
        space = self._user_overlaid_space(DEFAULT_TUNING_SPACES[stage])
        largest = 0
        for mbs in sorted(tuning_micro_batch_sizes):
            cfg = self._base_config(space, mbs)
            if max_train_batch_size_per_gpu and not tuning_micro_batch_sizes_overwritten:
                cfg[GRADIENT_ACCUMULATION_STEPS] = max(1, int(max_train_batch_size_per_gpu) // mbs)
            exp, val = self.run_ds_config(cfg, canonical_name(cfg, tuning_keys=[], prefix=tuning_space_name))
            self.update_records(tuning_space_name, exp, val, 1)
            if val is None:
                break
            largest = mbs
        return largest

    def _sweep_micro_batch(self, space, space_name, start_mbs, cap):
        """Doubling sweep; returns (largest runnable mbs, best mbs, best metric)."""
        mbs, largest, best_mbs, best_val = max(1, start_mbs), 0, 0, None
        stage = space.get(ZERO_OPTIMIZATION, {}).get(ZERO_OPTIMIZATION_STAGE, 0)
        budget = self.get_gpu_memory_info() - self.get_instantiation_memory_required_per_gpu(stage)
        act = self.get_activation_memory_per_gpu()
        while mbs <= cap:
            if act > 0 and mbs * act > budget and largest > 0:
                break  # the memory model says this cannot fit: do not spend an experiment on it
            cfg = self._base_config(space, mbs)
            name = canonical_name(cfg, tuning_keys=[], prefix=space_name)
            exp, val = self.run_ds_config(cfg, name)
            self.update_records(space_name, exp, val, 1)
            if val is None:
                break
            largest = mbs
            if best_val is None or (val < best_val if self.metric() == AUTOTUNING_METRIC_LATENCY else val > best_val):
                best_mbs, best_val = mbs, val
            mbs *= 2
        return largest, best_mbs, best_val

    def tune_space(self, tuning_space, prev_max_mbs=0, prev_best_mbs=0, prev_best_metric_val=0):
        stage = tuning_space.get(ZERO_OPTIMIZATION, {}).get(ZERO_OPTIMIZATION_STAGE, 0)
        space_name = f"z{stage}"
        user_mbs = self.user_config.get(TRAIN_MICRO_BATCH_SIZE_PER_GPU)
        start = self.min_train_micro_batch_size_per_gpu()
        if isinstance(user_mbs, int) and user_mbs > 0 and self.fast_enabled() is False:
            start = max(start, 1)
        largest, best_mbs, best_val = self._sweep_micro_batch(tuning_space, space_name, start, self.max_train_micro_batch_size_per_gpu())
        if largest == 0:
            logger.info(f"No runnable micro batch size found for ZeRO stage {stage}")
            return 0, 0, 0
        if self.fast_enabled() or not get_tuning_keys(tuning_space):
            return largest, best_mbs, best_val
        # full search: the stage's knobs x micro-batch candidates around the sweep's best point
        lo = max(self.min_train_micro_batch_size_per_gpu(), best_mbs // 2)
        mbs_list = self.get_tuning_micro_batch_size_list(lo, largest, self.num_tuning_micro_batch_sizes())
        space = copy.deepcopy(tuning_space)
        space[TRAIN_MICRO_BATCH_SIZE_PER_GPU] = mbs_list
        exps = self._generate_experiments(space, space_name)
        kind = self.autotuning_config.tuner_type
        if kind == AUTOTUNING_TUNER_MODELBASED:
            t = ModelBasedTuner(exps, self.rm, self.metric(), space)
        elif kind == AUTOTUNING_TUNER_RANDOM:
            t = RandomTuner(exps, self.rm, self.metric())
        else:
            t = GridSearchTuner(exps, self.rm, self.metric())
        sample = max(1, len(self.rm.nodes) * self.rm.nodes[0].max_slots // max(1, self.exp_num_gpus * self.exp_num_nodes)) if self.rm.nodes else 1
        n_run = t.tune(sample_size=sample, n_trials=self.autotuning_config.tuner_num_trials, early_stopping=self.autotuning_config.tuner_early_stopping)
        if t.best_exp is not None:
            self.update_records(space_name, t.best_exp, t.best_metric_val, n_run)
            if self._better(t.best_metric_val, best_val) or best_val is None:
                best_val = t.best_metric_val
                best_mbs = t.best_exp[DS_CONFIG][TRAIN_MICRO_BATCH_SIZE_PER_GPU]
        return largest, best_mbs, best_val

    def _generate_experiments(self, tuning_space, space_name):
        gas = self.get_gas_from_user_config()
        keys = get_tuning_keys(tuning_space)
        configs = prune_configs(get_all_configs(tuning_space, ignore_keys=["optimizer"]))
        exps = []
        for c in configs:
            cfg = copy.deepcopy(self.user_config)
            replace_dict(cfg, c)
            mbs = cfg[TRAIN_MICRO_BATCH_SIZE_PER_GPU]
            cfg[GRADIENT_ACCUMULATION_STEPS] = gas
            cfg[TRAIN_BATCH_SIZE] = mbs * gas * self.exp_num_gpus * self.exp_num_nodes // self.mp_size()
            exps.append(self._exp(canonical_name(cfg, keys, prefix=space_name), cfg))
        return exps

    # ------------------------------------------------------------------ driver
    def tune(self):
        self.start_time = time.time()
        if self.fast_enabled():
            logger.info("Fast mode is enabled. Tuning micro batch size only.")
        self.model_info = self.model_info_profile_run()
        if not self.model_info:
            return
        logger.info(f"The model has {number_to_string(self.get_model_num_params())} parameters.")
        self.gpu_mem = self.get_gpu_memory_info()
        logger.info(f"Memory per GPU in the system is {memory_to_string(self.gpu_mem, postfix='B')}.")
        self.activation_mem = self.get_activation_memory_per_gpu()
        logger.info(f"The model requires at least {memory_to_string(self.activation_mem, postfix='B')} activation memory for micro batch size 1.")
        stage = self.user_config.get(ZERO_OPTIMIZATION, {}).get(ZERO_OPTIMIZATION_STAGE, "all")
        allowed = stage if isinstance(stage, list) else [stage]
        logger.info(f"User-defined zero stages are {allowed}.")
        max_mbs, mbs, metric_val = 0, 0, None
        for s in (0, 1, 2, 3):
            if "all" not in allowed and s not in allowed:
                continue
            need = self.get_instantiation_memory_required_per_gpu(s) + self.activation_mem
            if self.gpu_mem <= need:
                logger.info(f"The model is not runnable with ZeRO stage {s} (which requires at least {memory_to_string(need, postfix='B')} memory with mbs = 1)")
                continue
            logger.info(f"The model might be runnable with ZeRO {s} (which requires at least {memory_to_string(need, postfix='B')} memory with mbs = 1), "
                        f"adding DEFAULT_TUNING_SPACE_ZERO_{s} to the global tuning space")
            space = self._user_overlaid_space(DEFAULT_TUNING_SPACES[s])
            nmax, nmbs, nval = self.tune_space(space, prev_max_mbs=max_mbs, prev_best_mbs=mbs, prev_best_metric_val=metric_val)
            if nval is not None and (metric_val is None or self._better(nval, metric_val) or nmbs > mbs):
                if metric_val is None or self._better(nval, metric_val):
                    metric_val = nval
                max_mbs, mbs = max(max_mbs, nmax), max(mbs, nmbs)
            elif nval is not None and metric_val is not None and s >= 1 and not self._better(nval, metric_val) and nmax <= max_mbs:
                logger.info(f"ZeRO stage {s} neither fits a larger micro batch nor improves {self.metric()}: stopping the stage search")
                break

    def _user_overlaid_space(self, default_space):
        """User-given zero_optimization values (scalars or lists) replace the defaults of the stage's tuning space."""
        space = copy.deepcopy(default_space)
        user_zero = copy.deepcopy(self.user_config.get(ZERO_OPTIMIZATION, {}))
        user_zero.pop(ZERO_OPTIMIZATION_STAGE, None)
        space[ZERO_OPTIMIZATION].update(user_zero)
        return space

    # ------------------------------------------------------------------ reporting
    def print_tuning_results(self):
        best = self.get_best_space_records()
        tab = []
        for key, (exp, val, num_exps) in best.items():
            if key == "global":
                continue
            tab.append([key, num_exps, val, exp["name"] if exp else ""])
        if "global" in best:
            exp, val, num_exps = best["global"]
            tab.append(["global", num_exps, val, exp["name"] if exp else ""])
        header = ["tuning_space", "num_experiments", "best_metric_val", "best_exp_name"]
        print(tabulate(tab, headers=header, tablefmt="pipe") if tabulate else "\n".join(str(r) for r in [header] + tab))
        if "global" in best:
            exp, val, total = best["global"]
            elapsed = datetime.timedelta(seconds=(time.time() - (self.start_time or time.time())))
            print(f"Tuning completed in {elapsed}. Total number of experiments: {total}.")
            if exp:
                print(f"{exp['name']} is the optimal setup after tuning. The exp result is at {exp.get('result_dir', self.results_dir)}.")
        with open(os.path.join(self.results_dir, "summary.txt"), "w") as fd:
            fd.write(json.dumps([[k, v[2], v[1], v[0]["name"] if v[0] else ""] for k, v in best.items()]))

    def write_optimal_config(self):
        best = self.get_best_space_records()
        if "global" not in best or best["global"][0] is None:
            return
        exp, _, _ = best["global"]
        ds_config = copy.deepcopy(exp[DS_CONFIG])
        ds_config.pop(AUTOTUNING, None)
        ds_config_path = os.path.join(self.results_dir, "ds_config_optimal.json")
        with open(ds_config_path, "w") as fd:
            json.dump(ds_config, fd)
        user_args = list(self.args.user_args)
        for flag in ("--deepspeed_config", "--deepspeed"):
            if flag in user_args and user_args.index(flag) + 1 < len(user_args):
                user_args[user_args.index(flag) + 1] = ds_config_path
        for key, arg in (self.autotuning_config.arg_mappings or {}).items():
            from .utils import get_val_by_key
            val = get_val_by_key(ds_config, key.strip())
            if val is not None and arg.strip() in user_args:
                user_args[user_args.index(arg.strip()) + 1] = str(val)
        cmd = [sys.executable, "-m", "deepspeedsyclsupport_b200.launcher.runner"]
        if getattr(self.args, "hostfile", None) and os.path.isfile(self.args.hostfile):
            cmd += ["--hostfile", self.args.hostfile]
        cmd += [self.args.user_script] + user_args
        with open(os.path.join(self.results_dir, "cmd_optimal.txt"), "w") as fd:
            fd.write(" ".join(cmd) + "\n")
        self.optimal_cmd = cmd
        self.optimal_ds_config = ds_config
        logger.info(f"Wrote the optimal DeepSpeed configuration found by autotuning to {ds_config_path}, and the corresponding command to "
                    f"{os.path.join(self.results_dir, 'cmd_optimal.txt')}")

    def run_after_tuning(self):
        """Launches the training with the optimal DeepSpeed configuration found through the autotuning process."""
        if self.optimal_cmd:
            result = subprocess.Popen(self.optimal_cmd)
            result.wait()
            logger.info(f"Done running with the optimal DeepSpeed configuration using {self.optimal_cmd}")
        else:
            logger.info("No optimal DeepSpeed configuration found by autotuning.")
