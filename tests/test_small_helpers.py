"""Helpers that have no other direct test: coalesced reduce-scatter, tuner feature maps, data-sampling utilities, CSV monitor,
MoE token mappings, the KV-split cost model, skinny-GEMM routing on CPU."""
import csv
import os

import numpy as np
import torch

from common import run_distributed


def _rsc_body():
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.runtime.comm.coalesced_collectives import reduce_scatter_coalesced
    r, W = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    base = [torch.randn(7), torch.randn(4, 3), torch.randn(1), torch.randn(10)]  # 7 and 1 leave short / empty chunks on the last ranks
    out = reduce_scatter_coalesced([b.clone() * (r + 1) for b in base])
    mean_factor = sum(range(1, W + 1)) / W
    for b, o in zip(base, out):
        c = -(-b.numel() // W)
        assert torch.allclose(o, (b.reshape(-1) * mean_factor)[r * c:r * c + c], atol=1e-6)
    even = torch.arange(6.0) * (r + 1)  # one tensor that divides evenly: reduced (and averaged) in place
    part = reduce_scatter_coalesced([even])[0]
    assert torch.allclose(part, (torch.arange(6.0) * mean_factor)[r * 2:r * 2 + 2])


def test_reduce_scatter_coalesced_three_ranks():
    run_distributed(_rsc_body, world_size=3, backend="gloo")


def test_tuner_feature_maps_round_trip():
    from deepspeedsyclsupport_b200.autotuning.tuner.utils import dict_to_dims, feature_to_index, gen_combinations, index_to_feature
    space = {"zero": {"stage": [1, 2, 3]}, "mbs": [1, 2, 4, 8], "fixed": 7, "flag": [True, False]}
    dims = dict_to_dims(space)
    assert dims == [3, 4, 1, 2]
    total = int(np.prod(dims))
    seen = set()
    for p in range(total):
        f = index_to_feature(p, dims)
        assert all(0 <= d < r for d, r in zip(f, dims)) and feature_to_index(f, dims) == p
        seen.add(tuple(f))
    assert len(seen) == total
    combos = list(gen_combinations(space))
    assert len(combos) == total and combos[0] == {"zero": {"stage": 1}, "mbs": 1, "fixed": 7, "flag": True}


def test_data_sampling_utils():
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_sampling.utils import find_fit_int_dtype, split_dataset, split_index
    assert find_fit_int_dtype(0, 255) == np.uint8 and find_fit_int_dtype(0, 256) == np.uint16 and find_fit_int_dtype(0, 2**40) == np.uint64
    assert find_fit_int_dtype(-1, 127) == np.int8 and find_fit_int_dtype(-129, 0) == np.int16 and find_fit_int_dtype(-1, 2**40) == np.int64
    assert split_index(0, 10, 3) == [(0, 4), (4, 8), (8, 10)] and split_index(3, 5, 4) == [(3, 4), (4, 5)]
    workers, threads = split_dataset(list(range(10)), 2, 1, 2)
    assert workers == [(0, 5), (5, 10)] and threads == [(5, 8), (8, 10)]


def test_csv_monitor_writes_one_file_per_event(tmp_path):
    from types import SimpleNamespace
    from deepspeedsyclsupport_b200.monitor.csv_monitor import csvMonitor
    mon = csvMonitor(SimpleNamespace(enabled=True, output_path=str(tmp_path), job_name="job"))
    mon.write_events([("Train/Samples/loss", 1.5, 10), ("Train/lr", 0.1, 10)])
    mon.write_events([("Train/Samples/loss", 1.25, 20)])
    rows = list(csv.reader(open(os.path.join(tmp_path, "job", "Train_Samples_loss.csv"))))
    assert rows == [["step", "loss"], ["10", "1.5"], ["20", "1.25"]]
    assert list(csv.reader(open(os.path.join(tmp_path, "job", "Train_lr.csv")))) == [["step", "lr"], ["10", "0.1"]]
    off = csvMonitor(SimpleNamespace(enabled=False, output_path=str(tmp_path), job_name="off"))
    off.write_events([("x", 1, 1)])
    assert not os.path.exists(os.path.join(tmp_path, "off"))


def test_moe_token_mappings_are_identity_without_tensor_parallelism():
    from deepspeedsyclsupport_b200.moe.mappings import drop_tokens, gather_tokens
    x = torch.randn(4, 6, requires_grad=True)
    assert gather_tokens(x, 0) is x and drop_tokens(x, 1) is x


def test_kv_split_cost_model():
    from deepspeedsyclsupport_b200.inference.v2.kernels import _pick_splits
    assert _pick_splits(2048, 8, 128, 129, 16) == 1            # a prefill chunk already has 129 CTAs: fp32 partials would only add traffic
    assert _pick_splits(64, 8, 128, 64, 34, page_us=0.45) == 2  # 64 decode atoms: two KV slices fill 128 of 148 SMs
    assert _pick_splits(8, 8, 128, 8, 257, page_us=0.45) == 16  # 8 sequences x 32K context: as many slices as allowed
    assert _pick_splits(1, 32, 128, 8, 2) <= 2                  # short context: nothing to split


def test_gemm_cpu_reference_path():
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    a, b = torch.randn(5, 16).bfloat16(), torch.randn(7, 16).bfloat16()
    out = gemm(a, b)
    assert out.shape == (5, 7) and torch.allclose(out.float(), a.float() @ b.float().t(), atol=0.1, rtol=0.05)


def test_alibi_slopes_are_read_off_the_tensor_and_arena_eligibility():
    from types import SimpleNamespace
    from deepspeedsyclsupport_b200.ops.transformer.inference.kv_arena import alibi_slopes_of, eligible
    heads, T = 4, 9
    slopes = 2.0 ** (-8.0 * torch.arange(1, heads + 1) / heads)
    alibi = (slopes[:, None] * torch.arange(T)[None, :]).view(heads, 1, T)  # [batch * heads, 1, T] as HF BLOOM builds it (batch 1)
    assert torch.allclose(alibi_slopes_of(alibi, 1, heads), slopes)
    assert alibi_slopes_of(alibi[..., :1], 1, heads) is None  # one key position: the slope cannot be recovered
    cfg = SimpleNamespace(triangular_masking=True, local_attention=False)
    x = torch.zeros(1, 3, 8, dtype=torch.bfloat16)
    assert not eligible(x, 64, heads, heads, cfg, None, None, None)  # CPU tensors never take the arena kernels


def test_lazy_llama_state_names_and_shapes():
    import bench_workloads
    cfg = dict(vocab_size=64, hidden_size=32, intermediate_size=48, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2)
    items = dict(bench_workloads._LazyLlamaState(cfg, "cpu", torch.float32).items())
    assert items["model.embed_tokens.weight"].shape == (64, 32) and items["lm_head.weight"].shape == (64, 32)
    assert items["model.layers.1.self_attn.k_proj.weight"].shape == (16, 32) and items["model.layers.0.mlp.down_proj.weight"].shape == (32, 48)
    assert len(items) == 3 + 2 * 9 and abs(float(items["model.norm.weight"].mean()) - 1.0) < 0.05


def test_fastgen_decode_runs_eagerly_on_cpu():
    """forward_graphed is the engine's entry point; without a CUDA device pure decode steps must simply run the eager forward."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_inference_v2 import _dense_llama_logits, _hf_llama_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    cfg = dict(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
               rms_norm_eps=1e-5)
    sd = _hf_llama_state(cfg)
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64, "max_ragged_sequence_count": 4, "max_context": 128,
                                                      "memory_config": {"mode": "allocate", "size": 32}})
    eng = build_engine_from_state_dict("llama", cfg, sd, ecfg, dtype=torch.float32)
    hist = torch.randint(0, 128, (11, ))
    eng.put([7], [hist])
    for _ in range(4):
        nxt = torch.randint(0, 128, (1, ))
        out = eng.put([7], [nxt])
        hist = torch.cat([hist, nxt])
        ref = _dense_llama_logits(sd, cfg, hist)[-1]
        assert (out[0].float() - ref).abs().max() < 5e-3 * max(1.0, ref.abs().max().item())
    assert not eng._model._graphs


def _obj_body():
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.runtime.pipe import p2p
    small = {"step": 3, "shape": (2, 5), "t": torch.arange(6).view(2, 3)}
    big = {"blob": torch.arange(5000, dtype=torch.float32), "names": ["x" * 50] * 40}
    if dist.get_rank() == 0:
        p2p.send_obj(small, 1)
        p2p.send_obj(big, 1)
    else:
        got = p2p.recv_obj(0)
        assert got["step"] == 3 and got["shape"] == (2, 5) and torch.equal(got["t"], small["t"])
        got = p2p.recv_obj(0)
        assert torch.equal(got["blob"], big["blob"]) and got["names"] == big["names"]


def test_pipe_send_recv_python_objects():
    run_distributed(_obj_body, world_size=2, backend="gloo")


def test_index_tuners_hand_out_every_experiment_once():
    import random
    from deepspeedsyclsupport_b200.autotuning.tuner.index_based_tuner import GridSearchTuner, RandomTuner
    exps = [{"name": f"e{i}", "ds_config": {}} for i in range(7)]
    grid = GridSearchTuner(list(exps), None, "throughput")
    assert [e["name"] for e in grid.next_batch(3)] == ["e0", "e1", "e2"] and [e["name"] for e in grid.next_batch(10)] == ["e3", "e4", "e5", "e6"]
    assert grid.next_batch(2) == []
    random.seed(0)
    rnd = RandomTuner(list(exps), None, "throughput")
    seen = []
    while True:
        batch = rnd.next_batch(2)
        if not batch:
            break
        seen += [e["name"] for e in batch]
    assert sorted(seen) == [f"e{i}" for i in range(7)] and seen != [f"e{i}" for i in range(7)]


def test_monitor_master_builds_only_enabled_backends(tmp_path):
    from types import SimpleNamespace
    from deepspeedsyclsupport_b200.monitor.monitor import MonitorMaster
    off = SimpleNamespace(enabled=False)
    cfg = SimpleNamespace(enabled=True, tensorboard=off, wandb=off,
                          csv_monitor=SimpleNamespace(enabled=True, output_path=str(tmp_path), job_name="run"))
    mm = MonitorMaster(cfg)
    assert mm.tb_monitor is None and mm.wandb_monitor is None and mm.csv_monitor is not None
    mm.write_events([("a/b", 1.0, 1)])
    assert os.path.exists(os.path.join(tmp_path, "run", "a_b.csv"))
    from deepspeedsyclsupport_b200.monitor import monitor as M
    assert M.csvMonitor.__name__ == "csvMonitor" and M.TensorBoardMonitor.__name__ == "TensorBoardMonitor"
