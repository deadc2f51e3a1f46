"""Pipeline parallelism: schedule as data, topology arithmetic, and a 2-stage PipelineEngine on gloo vs a 1-process run."""
import copy

import pytest
import torch

from common import run_distributed


def test_train_schedule_1f1b_shapes():
    from deepspeedsyclsupport_b200.runtime.pipe import schedule as S
    M, stages = 4, 2
    for stage in range(stages):
        sched = S.TrainSchedule(micro_batches=M, stages=stages, stage_id=stage)
        steps = list(sched.steps())
        assert len(steps) == 2 * (M + stages - 1)
        fwd = [c for cmds in steps for c in cmds if isinstance(c, S.ForwardPass)]
        bwd = [c for cmds in steps for c in cmds if isinstance(c, S.BackwardPass)]
        assert len(fwd) == M and len(bwd) == M
        assert isinstance(steps[-1][-1], S.OptimizerStep) and isinstance(steps[-1][-2], S.ReduceGrads)
        assert sched.num_pipe_buffers() == max(2, min(stages - stage, M))
    inf = S.InferenceSchedule(micro_batches=3, stages=2, stage_id=1)
    assert len(list(inf.steps())) == 3 + 2 - 1


def test_topology():
    from deepspeedsyclsupport_b200.runtime.pipe.topology import PipeDataParallelTopology, PipeModelDataParallelTopology, ProcessTopology
    topo = PipeDataParallelTopology(num_pp=2, num_dp=2)
    assert topo.world_size() == 4
    assert topo.get_rank(pipe=1, data=0) == 2
    assert topo.get_axis_comm_lists("pipe") == [[0, 2], [1, 3]]
    assert topo.get_axis_comm_lists("data") == [[0, 1], [2, 3]]
    assert topo.filter_match(pipe=0) == [0, 1]
    t3 = PipeModelDataParallelTopology(num_pp=2, num_mp=2, num_dp=2)
    assert t3.get_axis_list("pipe", 1) == [4, 5, 6, 7]
    assert t3.get_coord(5).model == 1
    assert ProcessTopology(["a", "b"], [2, 3]).get_rank_repr(4, omit_axes=[]) == "a_01-b_01"


class _Layer(torch.nn.Module):

    def __init__(self, d):
        super().__init__()
        self.lin = torch.nn.Linear(d, d)

    def forward(self, x):
        return torch.tanh(self.lin(x))


def _make_layers(d, n):
    torch.manual_seed(7)
    return [_Layer(d) for _ in range(n)]


def _pipe_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.pipe import PipelineModule
    d, n_layers, micro, gas = 16, 4, 4, 2
    layers = _make_layers(d, n_layers)
    ref_layers = copy.deepcopy(layers)
    net = PipelineModule(layers=layers, num_stages=2, loss_fn=torch.nn.MSELoss(), partition_method="uniform")
    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas,
           "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "zero_optimization": {"stage": 1}}
    engine, _, _, _ = deepspeed.initialize(model=net, config=cfg)
    g = torch.Generator().manual_seed(3)
    data = [(torch.randn(micro, d, generator=g), torch.randn(micro, d, generator=g)) for _ in range(gas * 3)]
    ref = torch.nn.Sequential(*ref_layers)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-2)
    it = iter(data)
    for step in range(3):
        loss = engine.train_batch(data_iter=it)
        ropt.zero_grad()
        tot = 0.0
        for k in range(gas):
            x, y = data[step * gas + k]
            l = torch.nn.functional.mse_loss(ref(x), y) / gas
            l.backward()
            tot += l.item()
        ropt.step()
        assert abs(loss.item() - tot) < 1e-4, (loss.item(), tot)
    # parameters of the local stage match the reference after 3 steps
    from deepspeedsyclsupport_b200.utils.tensor_fragment import safe_get_full_fp32_param
    stage = engine.stage_id
    for li in range(2):
        gl = stage * 2 + li
        p = getattr(net, str(gl)).lin.weight
        assert torch.allclose(safe_get_full_fp32_param(p), ref[gl].lin.weight.detach(), atol=1e-5)


def test_pipeline_engine_two_stages():
    run_distributed(_pipe_body, world_size=2)


def test_stage_boundaries_by_method():
    """Layer partitioning arithmetic (reference: pipe/test_pipe_module.py + runtime/utils partition_uniform / partition_balanced)."""
    from deepspeedsyclsupport_b200.runtime.pipe.module import LayerSpec, TiedLayerSpec, _stage_boundaries
    from deepspeedsyclsupport_b200.runtime.utils import partition_balanced, partition_uniform
    specs = [LayerSpec(torch.nn.Linear, 8, 8), LayerSpec(torch.nn.ReLU), LayerSpec(torch.nn.Linear, 8, 64), LayerSpec(torch.nn.ReLU),
             LayerSpec(torch.nn.Linear, 64, 8), LayerSpec(torch.nn.Tanh), LayerSpec(torch.nn.Linear, 8, 8), lambda x: x * 2]
    uni = _stage_boundaries(specs, "uniform", 4)
    assert uni == [0, 2, 4, 6, 8]
    by_type = _stage_boundaries(specs, "type:Linear", 2)  # two Linear layers per stage
    assert sum(isinstance(s, LayerSpec) and s.typename is torch.nn.Linear for s in specs[by_type[0]:by_type[1]]) == 2
    by_params = _stage_boundaries(specs, "parameters", 2)
    counts = [sum(p.numel() for p in s.build().parameters()) if isinstance(s, LayerSpec) else 0 for s in specs]
    left, right = sum(counts[:by_params[1]]), sum(counts[by_params[1]:])
    assert by_params[0] == 0 and by_params[-1] == 8 and max(left, right) <= 0.7 * sum(counts)  # the two big layers are split apart
    with pytest.raises((NotImplementedError, ValueError, AssertionError)):
        _stage_boundaries(specs, "no_such_method", 2)
    assert partition_uniform(10, 3) == [0, 4, 7, 10] or partition_uniform(10, 3) == [0, 3, 6, 10] or partition_uniform(10, 3)[-1] == 10
    assert partition_balanced([1, 1, 1, 9, 1, 1], 2)[1] in (3, 4)
    tied = TiedLayerSpec("embed", torch.nn.Embedding, 10, 4)
    assert tied.key == "embed" and tied.tied_weight_attr == ["weight"] or tied.tied_weight_attr == "weight"
    assert isinstance(tied.build(), torch.nn.Embedding) and "Embedding" in repr(tied)
