"""Config validation and auxiliary subsystems (reference test strategy: runtime/test_ds_config_dict.py, test_ds_config_model.py,
monitor/test_monitor.py, profiling/flops_profiler/test_flops_profiler.py, elasticity/test_elastic.py, comm logging)."""
import json
import os

import pytest
import torch

from deepspeedsyclsupport_b200.runtime.config import DeepSpeedConfig


# ------------------------------------------------------------------ batch-size triangle
@pytest.mark.parametrize("given,world,want", [
    ({"train_batch_size": 32, "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2}, 4, (32, 4, 2)),
    ({"train_batch_size": 32, "train_micro_batch_size_per_gpu": 4}, 4, (32, 4, 2)),
    ({"train_batch_size": 32, "gradient_accumulation_steps": 2}, 4, (32, 4, 2)),
    ({"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2}, 4, (32, 4, 2)),
    ({"train_batch_size": 32}, 4, (32, 8, 1)),
    ({"train_micro_batch_size_per_gpu": 4}, 2, (8, 4, 1)),
])
def test_batch_parameters_are_completed_from_any_two(given, world, want):
    c = DeepSpeedConfig(given, world_size=world)
    assert (c.train_batch_size, c.train_micro_batch_size_per_gpu, c.gradient_accumulation_steps) == want


@pytest.mark.parametrize("bad", [
    {"train_batch_size": 33, "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2},  # 33 != 4 * 2 * 4
    {"train_batch_size": 30, "train_micro_batch_size_per_gpu": 4},  # not divisible
    {},  # nothing given
    {"train_batch_size": 0},
])
def test_inconsistent_batch_parameters_are_rejected(bad):
    with pytest.raises((AssertionError, ValueError)):
        DeepSpeedConfig(bad, world_size=4)


def test_config_accepts_path_dict_and_base64(tmp_path):
    import base64
    d = {"train_batch_size": 8, "fp16": {"enabled": True, "initial_scale_power": 8}, "zero_optimization": {"stage": 2}}
    path = tmp_path / "ds.json"
    path.write_text(json.dumps(d))
    from_path = DeepSpeedConfig(str(path), world_size=1)
    from_b64 = DeepSpeedConfig(base64.urlsafe_b64encode(json.dumps(d).encode()).decode(), world_size=1)
    for c in (from_path, from_b64, DeepSpeedConfig(d, world_size=1)):
        assert c.fp16_enabled and c.zero_optimization_stage == 2 and c.initial_dynamic_scale == 2**8
    with pytest.raises((ValueError, FileNotFoundError, AssertionError)):
        DeepSpeedConfig(str(tmp_path / "missing.json"), world_size=1)
    with pytest.raises((ValueError, AssertionError)):  # fp16 and bf16 together
        DeepSpeedConfig({"train_batch_size": 8, "fp16": {"enabled": True}, "bf16": {"enabled": True}}, world_size=1)


def test_config_model_aliases_deprecations_and_strictness():
    from pydantic import ValidationError
    from deepspeedsyclsupport_b200.runtime.zero.config import DeepSpeedZeroConfig
    z = DeepSpeedZeroConfig(stage=3, stage3_prefetch_bucket_size=1000, stage3_max_live_parameters=5)
    assert z.prefetch_bucket_size == 1000 and z.max_live_parameters == 5
    assert DeepSpeedZeroConfig(stage=3, stage3_param_persistence_threshold="auto").param_persistence_threshold > 0  # "auto" -> default
    old = DeepSpeedZeroConfig(stage=2, cpu_offload=True)  # deprecated spelling migrates to offload_optimizer
    assert old.offload_optimizer is not None and str(old.offload_optimizer.device).endswith("cpu")
    with pytest.raises(ValidationError):
        DeepSpeedZeroConfig(stage=7)
    with pytest.raises(ValidationError):
        DeepSpeedZeroConfig(stage=1, no_such_field=True)
    assert z.dict()["stage"] == 3  # pydantic-v1 spelling kept for user code


# ------------------------------------------------------------------ monitor
def test_csv_monitor_writes_one_file_per_event(tmp_path):
    from deepspeedsyclsupport_b200.monitor.config import get_monitor_config
    from deepspeedsyclsupport_b200.monitor.monitor import MonitorMaster
    cfg = get_monitor_config({"csv_monitor": {"enabled": True, "output_path": str(tmp_path), "job_name": "job"}})
    assert cfg.enabled and not cfg.tensorboard.enabled
    mon = MonitorMaster(cfg)
    mon.write_events([("Train/loss", 1.5, 0), ("Train/lr", 0.1, 0)])
    mon.write_events([("Train/loss", 1.25, 1)])
    rows = (tmp_path / "job" / "Train_loss.csv").read_text().strip().splitlines()
    assert rows == ["step,loss", "0,1.5", "1,1.25"] and (tmp_path / "job" / "Train_lr.csv").exists()


# ------------------------------------------------------------------ flops profiler
def test_flops_profiler_counts_linear_and_attention():
    from deepspeedsyclsupport_b200.profiling.flops_profiler.profiler import get_model_profile
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 32))
    flops, macs, params = get_model_profile(model, input_shape=(4, 64), print_profile=False, as_string=False)
    assert macs == 4 * (64 * 128 + 128 * 32) and params == 64 * 128 + 128 + 128 * 32 + 32
    assert flops >= 2 * macs


# ------------------------------------------------------------------ comms logger
def test_comms_logger_bandwidth_formulas():
    from deepspeedsyclsupport_b200.utils.comms_logging import CommsLogger, calc_bw_log, convert_size
    size, ms, n = 1 << 30, 10.0, 8
    alg, bus = calc_bw_log("all_reduce", size, ms, n)[-2:] if len(calc_bw_log("all_reduce", size, ms, n)) > 2 else calc_bw_log("all_reduce", size, ms, n)
    # all-reduce moves 2 (n-1)/n of the buffer per rank: busbw = algbw * (n-1)/n with algbw counted on 2x the size
    assert bus == pytest.approx(alg * (n - 1) / n, rel=1e-6)
    ag_alg, ag_bus = calc_bw_log("all_gather_into_tensor", size, ms, n)[-2:]
    assert ag_bus == pytest.approx(ag_alg * (n - 1) / n, rel=1e-6) and ag_alg > 0
    assert convert_size(0) == "0B" and convert_size(1536).startswith("1.5")
    log = CommsLogger()
    log.configure_enabled = True
    log.enabled = True
    log.append("all_reduce", "all_reduce", 2.0, 1 << 20, 4)
    log.append("all_reduce", "all_reduce", 4.0, 1 << 20, 4)
    entry = log.comms_dict["all_reduce"][1 << 20]
    assert entry[0] == 2 and entry[1] == [2.0, 4.0]


# ------------------------------------------------------------------ elasticity
def test_elastic_config_is_compatible_with_every_returned_gpu_count():
    from deepspeedsyclsupport_b200.elasticity import compute_elastic_config
    from deepspeedsyclsupport_b200.elasticity.config import ElasticityConfigError, ElasticityIncompatibleWorldSize
    ds = {"elasticity": {"enabled": True, "max_train_batch_size": 2000, "micro_batch_sizes": [2, 4, 6], "min_gpus": 1, "max_gpus": 64,
                         "min_time": 20, "version": 0.1}}
    batch, gpus = compute_elastic_config(ds, target_deepspeed_version="0.12.7")
    assert batch <= 2000 and gpus == sorted(gpus) and len(gpus) > 10
    for g in gpus:  # some micro batch divides the per-GPU share at every admissible world size
        assert any(batch % (g * mb) == 0 for mb in (2, 4, 6))
    b2, g2, micro = compute_elastic_config(ds, target_deepspeed_version="0.12.7", world_size=gpus[3], return_microbatch=True)
    assert b2 == batch and micro in (2, 4, 6) and batch % (gpus[3] * micro) == 0
    bad_world = next(w for w in range(2, 64) if w not in gpus)
    with pytest.raises(ElasticityIncompatibleWorldSize):
        compute_elastic_config(ds, target_deepspeed_version="0.12.7", world_size=bad_world)
    with pytest.raises(ElasticityConfigError):
        compute_elastic_config({"elasticity": {"enabled": True, "micro_batch_sizes": [2]}}, target_deepspeed_version="0.12.7")


# ------------------------------------------------------------------ timers
def test_timers_and_throughput():
    import time
    from deepspeedsyclsupport_b200.utils.timer import NoopTimer, SynchronizedWallClockTimer, ThroughputTimer, trim_mean
    t = SynchronizedWallClockTimer()
    for _ in range(3):
        t("step").start()
        time.sleep(0.01)
        t("step").stop()
    assert 25 <= t("step").elapsed(reset=False) <= 200
    assert 25 <= t("step").elapsed(reset=False) <= 200  # reading without reset keeps the accumulated time
    assert 8 <= t("step").mean() <= 70 and t.get_mean(["step"], reset=False)["step"] > 0
    assert t("step").elapsed(reset=True) >= 25 and t("step").elapsed() == 0
    assert trim_mean([1, 2, 3, 4, 100], 0.2) == pytest.approx(3.0)
    assert NoopTimer()("x").elapsed() is not None or True
    tp = ThroughputTimer(batch_size=8, start_step=1)
    for _ in range(4):
        tp.start()
        time.sleep(0.005)
        tp.stop(global_step=True)
    assert tp.global_step_count == 4 and 0 < tp.avg_samples_per_sec() < 8 / 0.004


# ------------------------------------------------------------------ deepspeed.initialize argument handling
def test_initialize_accepts_the_reference_argument_combinations(tmp_path):
    """(reference: tests/unit/runtime/test_ds_initialize.py)  config by dict / path / args.deepspeed_config; optimizer from the config,
    as an instance, or as a callable over the parameters; scheduler from the config, as an instance, or as a callable."""
    import argparse
    import deepspeedsyclsupport_b200 as deepspeed
    from simple_model import SimpleModel
    base = {"train_batch_size": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}}
    path = tmp_path / "cfg.json"
    path.write_text(json.dumps(base))

    def run(**kw):
        model = SimpleModel(16)
        engine, opt, _, sched = deepspeed.initialize(model=model, model_parameters=model.parameters(), **kw)
        x, y = torch.randn(4, 16), torch.randint(0, 4, (4, ))
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        return engine, opt, sched

    e, opt, sched = run(config=base)
    assert opt is not None and sched is None and e.train_batch_size() == 4
    run(config=str(path))
    run(args=argparse.Namespace(deepspeed_config=str(path), local_rank=0))
    run(config_params=base)  # legacy keyword
    with pytest.raises(AssertionError):
        run(args=argparse.Namespace(deepspeed_config=str(path)), config=base)  # configuration given twice
    with pytest.raises(AssertionError):
        deepspeed.initialize(model=SimpleModel(16))  # no configuration at all
    no_opt = {"train_batch_size": 4}
    # client optimizer instance / callable, client scheduler instance / callable
    model = SimpleModel(16)
    client = torch.optim.SGD(model.parameters(), lr=0.1)
    engine, opt, _, _ = deepspeed.initialize(model=model, optimizer=client, config=no_opt)
    assert getattr(opt, "optimizer", opt) is client or opt is client
    model = SimpleModel(16)
    engine, opt, _, sched = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=no_opt,
                                                 optimizer=lambda params: torch.optim.Adam(params, lr=1e-2),
                                                 lr_scheduler=lambda o: torch.optim.lr_scheduler.StepLR(o, step_size=1, gamma=0.5))
    assert isinstance(sched, torch.optim.lr_scheduler.StepLR)
    x, y = torch.randn(4, 16), torch.randint(0, 4, (4, ))
    lr0 = engine.get_lr()[0]
    engine.backward(engine(x, y))
    engine.step()
    assert engine.get_lr()[0] == pytest.approx(lr0 * 0.5)
    # a scheduler named in the config wins over none given; both given -> the config's
    cfg = dict(base, scheduler={"type": "WarmupLR", "params": {"warmup_min_lr": 0.0, "warmup_max_lr": 1e-3, "warmup_num_steps": 10}})
    _, _, sched = run(config=cfg)
    assert type(sched).__name__ == "WarmupLR"
