"""ZeRO feature matrix on the host path (gloo): CPU offload, MiCS, ZeRO++ (qwZ / qgZ), fp16 dynamic loss scaling, frozen
parameters, zero.Init + GatheredParameters, tensor-fragment accessors, LR schedules driven from the config.
(reference tests: tests/unit/runtime/zero/*, tests/unit/runtime/half_precision/*, tests/unit/runtime/test_lr_schedulers.py)"""
import copy

import pytest
import torch

from common import run_distributed
from simple_model import SimpleFrozenModel, SimpleModel, random_batches
from test_zero_stages import _train


def test_cpu_offload_matches_plain_adamw_two_ranks():
    run_distributed(_train, world_size=2, args=(2, True, 1, 4, 1.0, "AdamW", {"offload_optimizer": {"device": "cpu", "pin_memory": False}}))


def test_zero3_cpu_offload_single_process():
    run_distributed(_train, world_size=1, args=(3, True, 2, 4, 1.0, "AdamW", {"offload_optimizer": {"device": "cpu"}}))


def _mics_body():
    _train(3, gas=1, steps=4, extra_zero={"mics_shard_size": 1, "stage3_param_persistence_threshold": 0})


def test_mics_shard_group_smaller_than_world():
    """mics_shard_size=1 at world 2: every rank holds full shards, gradients are all-reduced across the replica group."""
    run_distributed(_mics_body, world_size=2)


def _zeropp_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    torch.manual_seed(3)
    model = SimpleModel(hidden_dim=128, nlayers=2)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "zero_quantized_weights": True,
                                 "zero_quantized_gradients": True}}
    eng, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    losses = []
    x, y = random_batches(1, 4, 128, seed=7 + dist.get_rank())[0]  # one fixed batch per rank: the model must fit it
    for _ in range(12):
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(loss.item())
    assert all(torch.isfinite(torch.tensor(losses)))
    assert sum(losses[-3:]) < sum(losses[:3])  # fp8 block-scaled traffic still trains


def test_zeropp_quantized_weights_and_gradients_train():
    run_distributed(_zeropp_body, world_size=2)


def _fp16_body():
    import deepspeedsyclsupport_b200 as deepspeed
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=16, nlayers=2)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "fp16": {"enabled": True, "initial_scale_power": 40, "loss_scale_window": 2, "hysteresis": 1, "min_loss_scale": 1},
           "zero_optimization": {"stage": 1}}
    eng, opt, _, _ = deepspeed.initialize(model=model, config=cfg)
    assert next(model.parameters()).dtype == torch.float16
    scales, skipped = [], 0
    for x, y in random_batches(40, 4, 16, seed=5, dtype=torch.float16):
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        scales.append(opt.cur_scale)
        skipped += int(opt.overflow)
    assert skipped >= 1 and eng.skipped_steps == skipped  # 2^40 overflows: steps are skipped and the scale backs off
    assert min(scales) < 2.0**40 and scales[-1] >= min(scales)
    assert len(scales) - skipped >= 5  # and real steps are taken once the scale fits fp16


def test_fp16_dynamic_loss_scale_skips_and_recovers():
    run_distributed(_fp16_body, world_size=1)


def _frozen_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.utils.tensor_fragment import (safe_get_full_fp32_param, safe_get_full_grad, safe_get_full_optimizer_state,
                                                                  safe_set_full_fp32_param, safe_set_full_optimizer_state)
    torch.manual_seed(0)
    model = SimpleFrozenModel(hidden_dim=32)
    frozen_before = [p.detach().clone() for p in model.linears[0].parameters()]
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
    eng, _, _, _ = deepspeed.initialize(model=model, config=cfg, model_parameters=[p for p in model.parameters() if p.requires_grad])
    x, y = random_batches(1, 4, 32, seed=2)[0]
    loss = eng(x, y)
    eng.backward(loss)
    p = model.head.weight
    g = safe_get_full_grad(p)
    assert g is not None and g.shape == (4, 32) and g.abs().sum() > 0
    eng.step()
    with deepspeed.zero.GatheredParameters(list(model.linears[0].parameters())):
        for a, b in zip(frozen_before, model.linears[0].parameters()):
            assert torch.equal(a, b.detach())
    w = safe_get_full_fp32_param(p)
    assert w.shape == (4, 32) and safe_get_full_optimizer_state(p, "exp_avg").shape == (4, 32)
    safe_set_full_fp32_param(p, torch.full_like(w, 0.25))
    safe_set_full_optimizer_state(p, torch.zeros_like(w), "exp_avg")
    assert torch.allclose(safe_get_full_fp32_param(p), torch.full_like(w, 0.25))
    assert safe_get_full_optimizer_state(p, "exp_avg").abs().sum() == 0
    with deepspeed.zero.GatheredParameters([p]):
        assert torch.allclose(p.detach().float(), torch.full_like(w, 0.25))
    # the optimizer's own names for the same accessors + local (per-rank) fragments
    from deepspeedsyclsupport_b200 import comm as dist
    opt = eng.optimizer
    local = opt.get_local_fp32_param(p)
    n = torch.tensor([local.numel()])
    dist.all_reduce(n)
    assert n.item() == p.ds_numel and torch.all(local == 0.25)
    opt.set_local_hp_param(torch.full_like(local, 0.5), p)
    assert torch.allclose(opt.get_full_hp_param(p), torch.full_like(w, 0.5))
    assert opt.get_local_fp32_param(p, "exp_avg").numel() == local.numel() and opt.get_local_fp32_grad_for_param(p).numel() == local.numel()
    opt.set_lr(0.125)
    assert opt.get_lr()[0] == 0.125 and not opt.check_overflow()
    opt.override_loss_scale(4.0)
    assert opt.loss_scale == 4.0
    opt.destroy()


def test_frozen_params_and_fragment_accessors_two_ranks():
    run_distributed(_frozen_body, world_size=2)


def _init_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
    torch.manual_seed(11)
    with deepspeed.zero.Init(config_dict_or_path=cfg):
        model = SimpleModel(hidden_dim=32, nlayers=2)
    p = model.linears[0].weight
    assert hasattr(p, "ds_id") and p.ds_numel == 32 * 32 and p.numel() < p.ds_numel  # partitioned at construction
    with deepspeed.zero.GatheredParameters([p], modifier_rank=0):
        assert p.shape == (32, 32)
        if dist.get_rank() == 0:
            p.data.fill_(0.5)
    with deepspeed.zero.GatheredParameters([p]):
        assert torch.all(p.detach() == 0.5)  # rank 0's modification reached every shard
    eng, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    losses = []
    x, y = random_batches(1, 4, 32, seed=1)[0]
    for _ in range(6):
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]


def test_zero_init_partitions_at_construction_two_ranks():
    run_distributed(_init_body, world_size=2)


def test_lr_schedules_from_config():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.runtime import lr_schedules as S
    model = SimpleModel(hidden_dim=8, nlayers=1)
    opt = torch.optim.SGD(model.parameters(), lr=1.0)
    w = S.WarmupLR(opt, warmup_min_lr=0.0, warmup_max_lr=0.1, warmup_num_steps=10, warmup_type="linear")
    lrs = []
    for _ in range(12):
        w.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert abs(lrs[5] - 0.05) < 1e-6 and abs(lrs[-1] - 0.1) < 1e-9
    d = S.WarmupDecayLR(opt, total_num_steps=20, warmup_min_lr=0.0, warmup_max_lr=0.1, warmup_num_steps=10, warmup_type="linear")
    for _ in range(20):
        d.step()
    assert opt.param_groups[0]["lr"] < 0.011
    c = S.WarmupCosineLR(opt, total_num_steps=20, warmup_num_steps=5, warmup_min_ratio=0.0, cos_min_ratio=0.1)
    vals = []
    for _ in range(20):
        c.step()
        vals.append(opt.param_groups[0]["lr"])
    assert max(vals) <= 1.0 + 1e-6 and vals[-1] < vals[6]
    oc = S.OneCycle(opt, cycle_min_lr=0.01, cycle_max_lr=0.1, cycle_first_step_size=5)
    seq = []
    for _ in range(10):
        oc.step()
        seq.append(opt.param_groups[0]["lr"])
    assert abs(max(seq) - 0.1) < 1e-6 and seq[0] < seq[3] and seq[-1] < seq[4]
    rt = S.LRRangeTest(opt, lr_range_test_min_lr=0.001, lr_range_test_step_size=2, lr_range_test_step_rate=1.0)
    rt.step()
    rt.step()
    rt.step()
    assert opt.param_groups[0]["lr"] > 0.001


def _nvme_body(path):
    """ZeRO-Infinity optimizer tier: state streamed from files in sub-groups must reproduce the host-offload run exactly, and the
    slow paths (fragment accessors, state_dict round trip) must see / persist the same values."""
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state

    def run(offload):
        torch.manual_seed(0)
        model = SimpleModel(hidden_dim=64, nlayers=3)
        cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "sub_group_size": 3000, "offload_optimizer": offload},
               "aio": {"block_size": 262144, "queue_depth": 4, "thread_count": 1}}
        eng, _, _, _ = deepspeed.initialize(model=model, config=cfg)
        losses = []
        for i, (x, y) in enumerate(random_batches(5, 4, 64, seed=3)):
            loss = eng(x, y)
            eng.backward(loss)
            eng.step()
            losses.append(loss.item())
            if i == 2:  # a slow path in the middle of training: the region becomes resident and is written back by the next step
                p = model.linears[1].weight
                assert safe_get_full_optimizer_state(p, "exp_avg").abs().sum() > 0
        p = model.linears[0].weight
        return losses, safe_get_full_fp32_param(p).clone(), eng

    l_cpu, w_cpu, _ = run({"device": "cpu"})
    l_nvme, w_nvme, eng = run({"device": "nvme", "nvme_path": path})
    assert l_cpu == l_nvme and torch.equal(w_cpu, w_nvme)
    region = eng.optimizer.regions[0]
    assert getattr(region, "is_nvme", False) and len(region.bounds) > 1
    import os
    files = [f for _, _, fs in os.walk(path) for f in fs if f.endswith(".swp")]
    assert len(files) == 3 * len(region.bounds)  # master, exp_avg, exp_avg_sq per sub-group
    sd = eng.optimizer.state_dict()
    eng.optimizer.load_state_dict(sd)


def test_nvme_optimizer_state_matches_host_offload(tmp_path):
    run_distributed(_nvme_body, world_size=1, args=(str(tmp_path), ))


def _offload_param_body(path, device, world_check):
    """ZeRO-Infinity parameter tier (host memory or a memory-mapped file): same losses as plain ZeRO-3; units are only
    resident between fetch and release."""
    import deepspeedsyclsupport_b200 as deepspeed

    def run(zero_extra):
        torch.manual_seed(0)
        model = SimpleModel(hidden_dim=64, nlayers=3)
        cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
               "zero_optimization": dict({"stage": 3, "stage3_param_persistence_threshold": 0}, **zero_extra)}
        eng, _, _, _ = deepspeed.initialize(model=model, config=cfg)
        losses = []
        for x, y in random_batches(5, 4, 64, seed=3):
            loss = eng(x, y)
            eng.backward(loss)
            eng.step()
            losses.append(loss.item())
        return losses, eng, model

    base, _, _ = run({})
    op = {"device": device, "nvme_path": path} if device == "nvme" else {"device": "cpu"}
    got, eng, model = run({"offload_param": op, "offload_optimizer": {"device": "cpu"}})
    assert all(abs(a - b) < 1e-5 for a, b in zip(base, got)), (base, got)
    opt = eng.optimizer
    assert opt.offload_param and opt.lp_shard.device.type == "cpu" and not opt.keep_full
    assert all(u.full is None for u in opt.units if not u.persistent)  # released after the step
    if device == "nvme":
        import os
        assert os.path.getsize(opt._lp_file) == opt.total_shard * 4
    with deepspeed.zero.GatheredParameters([model.linears[0].weight]):
        assert model.linears[0].weight.shape == (64, 64)


@pytest.mark.parametrize("device,world", [("cpu", 1), ("cpu", 2), ("nvme", 1)])
def test_offload_param_matches_plain_zero3(tmp_path, device, world):
    run_distributed(_offload_param_body, world_size=world, args=(str(tmp_path), device, world))


def test_contiguous_memory_allocator_compacts_and_repoints():
    from deepspeedsyclsupport_b200.runtime.zero.contiguous_memory_allocator import ContiguousMemoryAllocator
    a = ContiguousMemoryAllocator(64, torch.float32, "cpu")
    ts = [a.allocate_tensor(16) for _ in range(4)]
    for i, t in enumerate(ts):
        t.fill_(i + 1)
    p = torch.nn.Parameter(torch.empty(0))
    a.assign_to_param(ts[3], p, 12, (3, 4))
    a.release_tensor(ts[0])
    a.release_tensor(ts[2])
    assert a.total_free == 32 and a._largest_contiguous() == 16 and sorted(a.contiguous_sizes.values()) == [16, 16]
    big = a.allocate_tensor(32)  # only possible after sliding the two live tensors together
    big.fill_(9)
    assert ts[1].eq(2).all() and ts[3].eq(4).all() and p.data.eq(4).all() and p.shape == (3, 4)
    assert ts[1].data_ptr() == a.buffer.data_ptr() and p.data_ptr() == ts[3].data_ptr() and a.max_allocated == 64
    a.release_tensor(ts[3])
    assert p.numel() == 0 and a.total_free == 16
    with pytest.raises(AssertionError):
        a.allocate_tensor(17)


def test_zero3_linear_function_matches_torch_linear():
    from deepspeedsyclsupport_b200.runtime.zero.linear import LinearModuleForZeroStage3, zero3_linear_wrap
    torch.manual_seed(0)
    x = torch.randn(2, 5, 8, requires_grad=True)
    ref = torch.nn.Linear(8, 6)
    mod = LinearModuleForZeroStage3(8, 6)
    mod.load_state_dict(ref.state_dict())
    (mod(x) ** 2).sum().backward()
    gx, gw, gb = x.grad.clone(), mod.weight.grad.clone(), mod.bias.grad.clone()
    x.grad = None
    (ref(x) ** 2).sum().backward()
    assert torch.allclose(gx, x.grad, atol=1e-5) and torch.allclose(gw, ref.weight.grad, atol=1e-5) and torch.allclose(gb, ref.bias.grad, atol=1e-5)
    assert torch.allclose(zero3_linear_wrap(x, ref.weight), torch.nn.functional.linear(x, ref.weight), atol=1e-6)


def test_mics_group_arithmetic_and_profiler_counters():
    from deepspeedsyclsupport_b200.runtime.zero.mics_utils import _generate_mics_config, _sizes_all_same, scale_tensors
    from deepspeedsyclsupport_b200.runtime.zero.partitioned_param_profiler import PartitionedParameterProfiler
    from deepspeedsyclsupport_b200.utils.timer import SynchronizedWallClockTimer
    cfg = _generate_mics_config(world_size=16, ndev_per_node=8, shard_size=4)
    assert cfg["shard_groups"][1] == [4, 5, 6, 7] and cfg["replicate_groups"][2] == [2, 6, 10, 14] and _sizes_all_same(cfg["shard_groups"])
    with pytest.raises(AssertionError):
        _generate_mics_config(world_size=10, ndev_per_node=8, shard_size=4)
    t = [torch.ones(3), torch.full((2, ), 4.0)]
    scale_tensors(t, 2)
    assert t[0].tolist() == [0.5] * 3 and t[1].tolist() == [2.0, 2.0]
    prof = PartitionedParameterProfiler(SynchronizedWallClockTimer())
    for n in (10, 20):
        prof.start_event("fetch")
        prof.stop_event("fetch", n)
    c = prof.event_counters["fetch"]
    assert (c.count, c.num_elem) == (2, 30)
    prof.reset_events()
    assert prof.event_counters["fetch"].count == 0


def _offload_body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.runtime.zero.parameter_offload import DeepSpeedZeRoOffload
    from deepspeedsyclsupport_b200.runtime.zero.partitioned_param_coordinator import ZeRoTraceMode
    from deepspeedsyclsupport_b200.runtime.zero.utils import assert_ints_same_as_other_ranks, get_lst_from_rank0
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.ReLU(), torch.nn.Linear(32, 8))
    x = torch.randn(4, 16)
    with torch.no_grad():
        ref = model(x)
    off = DeepSpeedZeRoOffload(model, param_persistence_threshold=0, dtype=torch.float32)
    # parameters are sharded: nothing gathered between forwards
    assert all(p.numel() == 0 or p.ds_status.name == "NOT_AVAILABLE" for p in model.parameters())
    coord = off.get_param_coordinator(training=False)
    assert coord.trace_mode == ZeRoTraceMode.RECORD
    with torch.no_grad():
        out = model(x)
        coord.reset_step()
        assert coord.is_complete_trace()
        out2 = model(x)
    assert torch.allclose(out, ref, atol=1e-6) and torch.allclose(out2, ref, atol=1e-6)
    coord.fetch_sub_module(model[0], forward=True)
    assert model[0].weight.shape == (32, 16)
    coord.release_sub_module(model[0], backward=False)
    off.partition_all_parameters()
    assert get_lst_from_rank0([dist.get_rank() + 5, 7]) == [5, 7]
    assert_ints_same_as_other_ranks([1, 2, 3])
    if dist.get_rank() == 1:
        with pytest.raises(RuntimeError):
            assert_ints_same_as_other_ranks([9])
    else:
        assert_ints_same_as_other_ranks([0])
    off.destroy()


def test_zero_offload_facade_runs_sharded_inference_two_ranks():
    from common import run_distributed
    run_distributed(_offload_body, world_size=2)
