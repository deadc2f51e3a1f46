"""Autotuner search logic with an injected experiment runner (no GPUs): the fake cluster OOMs above a stage-dependent
micro-batch size and has a known best configuration; reference tests: tests/unit/autotuning/test_autotuning.py."""
import argparse
import json
import os

import pytest


def _fake_runner(log):

    def run(exp, reservations, user_script, user_args):
        cfg = exp["ds_config"]
        at = cfg["autotuning"]
        os.makedirs(exp["result_dir"], exist_ok=True)
        open(os.path.join(exp["result_dir"], "stderr.log"), "w").close()
        if (at.get("model_info") or {}).get("profile"):
            os.makedirs(os.path.dirname(at["model_info_path"]), exist_ok=True)
            json.dump({"num_params": 1_000_000_000, "trainable_num_params": 1_000_000_000, "activation_mem_per_gpu": 4 * 1024**3},
                      open(at["model_info_path"], "w"))
            return
        stage = cfg.get("zero_optimization", {}).get("stage", 0)
        mbs = cfg["train_micro_batch_size_per_gpu"]
        log.append((stage, mbs, cfg.get("zero_optimization", {}).get("overlap_comm")))
        limit = {0: 4, 1: 8, 2: 16, 3: 16}[stage]
        if mbs > limit:
            with open(os.path.join(exp["result_dir"], "stderr.log"), "w") as f:
                f.write("RuntimeError: CUDA out of memory\n")
            return
        # throughput grows with mbs, stage 2 is the sweet spot, stage 3 pays a gather tax; overlap_comm helps a little
        tput = mbs * 100 * {0: 1.0, 1: 1.0, 2: 1.05, 3: 0.8}[stage] * (1.1 if cfg.get("zero_optimization", {}).get("overlap_comm") else 1.0)
        json.dump({"throughput": tput, "latency": 1000.0 / tput, "FLOPS_per_gpu": tput * 1e9}, open(at["metric_path"], "w"))

    return run


def _args(tmp, cfg):
    path = os.path.join(tmp, "ds_config.json")
    json.dump(cfg, open(path, "w"))
    return argparse.Namespace(user_script="train.py", user_args=["--deepspeed_config", path, "--foo", "1"], hostfile="/nonexistent", master_port=29500,
                              launcher_args="", num_nodes=-1, num_gpus=-1)


@pytest.mark.parametrize("fast", [True, False])
def test_autotuner_finds_best_stage_and_mbs(tmp_path, fast):
    from deepspeedsyclsupport_b200.autotuning import Autotuner
    tmp = str(tmp_path)
    cfg = {"train_micro_batch_size_per_gpu": 1, "bf16": {"enabled": True}, "zero_optimization": {"stage": "all"},
           "autotuning": {"enabled": True, "fast": fast, "results_dir": os.path.join(tmp, "results"), "exps_dir": os.path.join(tmp, "exps"),
                          "tuner_type": "gridsearch", "tuner_num_trials": 12, "tuner_early_stopping": 6, "max_train_micro_batch_size_per_gpu": 64}}
    log = []
    tuner = Autotuner(_args(tmp, cfg), {"localhost": list(range(8))}, run_fn=_fake_runner(log))
    tuner.tune()
    assert tuner.get_model_num_params() == 1_000_000_000
    tuner.print_tuning_results()
    tuner.write_optimal_config()
    best = json.load(open(os.path.join(tmp, "results", "ds_config_optimal.json")))
    assert best["zero_optimization"]["stage"] == 2 and best["train_micro_batch_size_per_gpu"] == 16
    assert best["train_batch_size"] == 16 * 8 and "autotuning" not in best
    if not fast:
        assert best["zero_optimization"]["overlap_comm"] is True
    cmd = open(os.path.join(tmp, "results", "cmd_optimal.txt")).read()
    assert "ds_config_optimal.json" in cmd and "train.py" in cmd
    # the doubling sweep stopped right after the first failing micro-batch of each stage
    assert max(m for s, m, _ in log if s == 0) == 8 and max(m for s, m, _ in log if s == 2) == 32


def test_tuning_space_helpers():
    from deepspeedsyclsupport_b200.autotuning.utils import canonical_name, get_all_configs, get_first_config, get_tuning_keys, prune_configs
    space = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False], "reduce_bucket_size": [1, 2, 3]}, "train_micro_batch_size_per_gpu": [4, 8]}
    assert sorted(get_tuning_keys(space)) == ["overlap_comm", "reduce_bucket_size", "train_micro_batch_size_per_gpu"]
    configs = get_all_configs(space)
    assert len(configs) == 12 and len(prune_configs(configs)) == 12
    assert get_first_config(space) == {"zero_optimization": {"stage": 2, "overlap_comm": True, "reduce_bucket_size": 1}, "train_micro_batch_size_per_gpu": 4}
    name = canonical_name({"zero_optimization": {"stage": 2, "overlap_comm": True}, "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 1},
                          ["overlap_comm"], prefix="z2")
    assert name.startswith("z2_") and "tmbspg4" in name and "gas1" in name


def test_resource_manager_packs_experiments(tmp_path):
    from deepspeedsyclsupport_b200.autotuning.scheduler import ResourceManager
    from deepspeedsyclsupport_b200.autotuning.utils import write_experiments
    import threading
    tmp = str(tmp_path)
    live, peak = [0], [0]
    lock = threading.Lock()

    def run(exp, reservations, user_script, user_args):
        import time
        with lock:
            live[0] += 1
            peak[0] = max(peak[0], live[0])
        time.sleep(0.3)
        os.makedirs(exp["result_dir"], exist_ok=True)
        json.dump({"throughput": float(exp["exp_id"])}, open(exp["ds_config"]["autotuning"]["metric_path"], "w"))
        with lock:
            live[0] -= 1

    args = argparse.Namespace(user_script="t.py", user_args=[], hostfile=None, master_port=29500, launcher_args="")
    rm = ResourceManager(args, ["localhost"], 8, os.path.join(tmp, "res"), os.path.join(tmp, "exps"), None, run_fn=run)
    exps = [{"name": f"e{i}", "ds_config": {"autotuning": {}}, "num_gpus": 4, "num_nodes": 1} for i in range(4)]
    rm.schedule_experiments(write_experiments(exps, rm.exps_dir))
    rm.run()
    best, val = rm.parse_results("throughput")
    assert best["name"] == "e3" and val == 3.0 and peak[0] == 2  # two 4-GPU experiments share the 8-GPU node
