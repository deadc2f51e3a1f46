"""Host-side tests of the smaller public helpers: overflow check, tensor-list norms, swap buffers, fragment records,
scheduler config helpers, packed quantized parameters, env patching."""
import math
import os
import types

import pytest
import torch

import deepspeedsyclsupport_b200 as deepspeed
from deepspeedsyclsupport_b200.runtime import utils as ru


def test_top_level_names():
    assert deepspeed.ADAM_OPTIMIZER == "adam" and deepspeed.LAMB_OPTIMIZER == "lamb"
    from deepspeedsyclsupport_b200.compression import compression_scheduler  # noqa: F401
    from deepspeedsyclsupport_b200.module_inject import HFBertLayerPolicy  # noqa: F401
    from deepspeedsyclsupport_b200.ops.transformer import (DeepSpeedInferenceConfig, DeepSpeedMoEInference,  # noqa: F401
                                                           DeepSpeedTransformerInference)
    from deepspeedsyclsupport_b200.runtime.pipe import p2p
    assert p2p.can_send_recv()


def test_check_overflow():
    ps = [torch.nn.Parameter(torch.randn(5)) for _ in range(3)]
    for p in ps:
        p.grad = torch.randn(5)
    chk = ru.CheckOverflow([ps])
    assert chk.check() is False
    ps[1].grad[2] = float("inf")
    assert chk.check() is True
    ps[1].grad[2] = float("nan")
    assert chk.has_overflow(ps) is True
    assert chk.check_using_norm([1.0, -1]) is True and chk.check_using_norm([1.0, 2.0]) is False
    assert ru.CheckOverflow._has_inf_or_nan(torch.tensor([1.0, float("nan")]))


def test_tensor_list_norm_and_clip():
    ts = [torch.randn(7), torch.randn(3, 4), torch.randn(2)]
    ref = torch.cat([t.flatten() for t in ts]).norm().item()
    assert math.isclose(ru.get_global_norm_of_tensors(ts), ref, rel_tol=1e-5)
    assert math.isclose(ru.get_global_norm_of_tensors(ts, norm_type=math.inf), max(t.abs().max().item() for t in ts), rel_tol=1e-6)
    assert ru.get_global_norm_of_tensors([torch.tensor([float("nan")])]) == -1
    before = ru.clip_tensors_by_global_norm(ts, max_norm=0.5)
    assert math.isclose(before, ref, rel_tol=1e-5)
    assert math.isclose(torch.cat([t.flatten() for t in ts]).norm().item(), 0.5, rel_tol=1e-4)


def test_align_tlinear_copy_version():
    out = ru.align_dense_tensors([torch.ones(5), torch.ones(6)], 8)
    assert sum(t.numel() for t in out) == 16 and out[-1].abs().sum() == 0
    assert ru.align_dense_tensors([torch.ones(8)], 8)[0].numel() == 8
    holder = types.SimpleNamespace(weight=torch.nn.Parameter(torch.randn(6, 4)), bias=torch.nn.Parameter(torch.randn(4)))
    x = torch.randn(3, 6)
    torch.testing.assert_close(ru.TLinear(holder)(x), x @ holder.weight + holder.bias)
    src = {"a": [torch.ones(2)], "b": 3}
    cp = ru.copy_to_device(src, "cpu", torch.is_tensor)
    assert cp["a"][0].data_ptr() != src["a"][0].data_ptr() and cp["b"] == 3
    assert ru.required_torch_version(min_version=1.8) and not ru.required_torch_version(max_version=1.0)
    calls = []
    ru.graph_process(False, lambda t: calls.append(t), 1)
    assert calls == [1] or torch.cuda.is_available()
    assert ru.get_inactive_params([torch.nn.Parameter(torch.ones(1))]) == []


def test_swap_buffers():
    from deepspeedsyclsupport_b200.runtime.swap_tensor.utils import SwapBuffer, SwapBufferManager, get_sized_buffers
    buf = SwapBuffer(torch.zeros(64))
    t = torch.arange(10.0)
    swap, comp = buf.insert_tensor(t, "/tmp/a.swp", 16)
    assert swap.numel() == 16 and torch.equal(comp, t)
    buf.allocate_tensor("/tmp/b.swp", 20, 32)
    assert buf.get_swap_paths() == ["/tmp/a.swp", "/tmp/b.swp"] and buf.get_num_elem() == 30
    assert buf.has_space(16) and not buf.has_space(17) and buf.get_swap_path(16) == "/tmp/b.swp"
    buf.reset()
    assert buf.get_swap_tensors() == [] and buf.has_space(64)
    mgr = SwapBufferManager(num_elems=32, count=3, dtype=torch.float32)
    a = mgr.allocate(16, 2, torch.float32)
    assert len(a) == 2 and a[0].numel() == 16 and mgr.allocate(16, 2, torch.float32) is None
    mgr.free(a)
    assert len(mgr.allocate_all(32, torch.float32)) == 3
    assert [v.numel() for v in get_sized_buffers([torch.zeros(8), torch.zeros(8)], [3, 8])] == [3, 8]


def test_fragment_mapping():
    from deepspeedsyclsupport_b200.utils.tensor_fragment import get_hp_fragment_mapping
    lp = torch.arange(12.0).bfloat16().reshape(3, 4)  # occupies [10, 22) of the flat group
    hp_partition = torch.zeros(8)  # this rank owns [16, 24)
    state = {"exp_avg": torch.zeros(8), "step": torch.tensor(1)}
    frag = get_hp_fragment_mapping(lp, 10, hp_partition, {}, {}, False, 0, 16, 8, state)
    assert (frag.lp_fragment_address.start, frag.lp_fragment_address.numel) == (6, 6)
    assert (frag.hp_fragment_address.start, frag.hp_fragment_address.numel) == (0, 6)
    frag.update_hp()
    torch.testing.assert_close(hp_partition[:6], torch.arange(6.0, 12.0))
    assert frag.get_optim_state_keys() == ["exp_avg"] and frag.get_hp_fragment("exp_avg").numel() == 6
    with pytest.raises(ValueError):
        frag.get_optim_state_fragment("exp_avg_sq")


def test_lr_config_helpers():
    import argparse
    from deepspeedsyclsupport_b200.runtime import lr_schedules as ls
    parser = ls.add_tuning_arguments(argparse.ArgumentParser())
    args = parser.parse_args(["--lr_schedule", "OneCycle", "--cycle_max_lr", "0.3"])
    cfg, err = ls.get_config_from_args(args)
    assert err is None and cfg["type"] == "OneCycle" and cfg["params"]["cycle_max_lr"] == 0.3
    assert ls.get_lr_from_config(cfg) == (0.3, "")
    assert ls.get_config_from_args(parser.parse_args([]))[0] is None
    assert ls.get_lr_from_config({"type": "Nope", "params": {}})[0] is None
    args = parser.parse_args(["--lr_schedule", "WarmupLR", "--warmup_max_lr", "0.01"])
    assert ls.get_lr_from_config(ls.get_config_from_args(args)[0])[0] == 0.01


def test_packed_quantized_param():
    import torch.nn.functional as F
    from deepspeedsyclsupport_b200.inference.quantization import utils as qu
    cfg = {"num_bits": 8, "group_size": 16, "group_dim": 1, "symmetric": False}
    lin = torch.nn.Linear(32, 8, bias=False)
    w = lin.weight.data.clone()
    qu._quantize_param(lin.weight, cfg)
    assert lin.weight.dtype == torch.uint8 and lin.weight.weight_quantized
    x = torch.randn(4, 32)
    y = qu.wrap_quantized_functional(F.linear)(x, lin.weight)
    torch.testing.assert_close(y, x @ w.t(), atol=5e-2, rtol=5e-2)
    sd = {"weight": torch.randn(8, 32)}
    keep = sd["weight"]
    seen = {}
    qu.wrap_load_from_state_dict(lambda m, s, *a: seen.update(dtype=s["weight"].dtype))(lin, sd, "", {}, True, [], [], [])
    assert seen["dtype"] == torch.uint8 and sd["weight"] is keep
    h = torch.tensor([1.4, -3.6], dtype=torch.float16)
    assert qu.tensor_round(h.clone()).tolist() == [1.0, -4.0] and qu.tensor_clamp(h.clone(), -1, 1).tolist() == [1.0, -1.0]


def test_env_patching(monkeypatch):
    from deepspeedsyclsupport_b200 import comm as dist
    for k, v in dict(OMPI_COMM_WORLD_RANK="3", OMPI_COMM_WORLD_LOCAL_RANK="1", OMPI_COMM_WORLD_SIZE="8",
                     OMPI_COMM_WORLD_LOCAL_SIZE="4", AZ_BATCH_MASTER_NODE="10.0.0.4:6000").items():
        monkeypatch.setenv(k, v)
    # the helpers write os.environ directly: register every key they touch with monkeypatch first, so that the
    # values are rolled back and cannot leak a fake rendezvous address into later tests
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        monkeypatch.setenv(k, "placeholder")
        del os.environ[k]
    dist.patch_aml_env_for_torch_nccl_backend(verbose=False)
    assert (os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"]) == ("3", "1", "8")
    assert os.environ["MASTER_ADDR"] == "10.0.0.4" and os.environ["MASTER_PORT"] == "6105"
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "5")
    dist.patch_aws_sm_env_for_torch_nccl_backend(verbose=False)
    assert os.environ["RANK"] == "5"


def test_mics_handle():
    from deepspeedsyclsupport_b200.runtime.zero.mics import MiCS_AllGatherCoalescedHandle
    from deepspeedsyclsupport_b200.runtime.zero.partition_parameters import ZeroParamStatus
    p = torch.nn.Parameter(torch.ones(1))
    p.ds_status = ZeroParamStatus.INFLIGHT
    waited = []
    h = MiCS_AllGatherCoalescedHandle(types.SimpleNamespace(wait=lambda: waited.append(1)), [p], [None], 2)
    h.wait(); h.wait()
    assert waited == [1] and p.ds_status == ZeroParamStatus.AVAILABLE


def test_env_patching_left_no_trace():
    assert os.environ.get("MASTER_ADDR") != "10.0.0.4" and os.environ.get("WORLD_SIZE") != "8"
