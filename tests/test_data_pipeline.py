"""Data-efficiency library: curriculum scheduler, mmap indexed dataset, analyzer -> sampler round trip, random-LTD
(reference tests: tests/unit/runtime/test_data_efficiency.py, tests/unit/ops/... random_ltd)."""
import numpy as np
import torch

from deepspeedsyclsupport_b200.runtime.data_pipeline.constants import *  # noqa: F401,F403


def test_curriculum_schedules():
    from deepspeedsyclsupport_b200.runtime.data_pipeline.curriculum_scheduler import CurriculumScheduler
    s = CurriculumScheduler({"min_difficulty": 8, "max_difficulty": 64, "schedule_type": "fixed_linear",
                             "schedule_config": {"total_curriculum_step": 100, "difficulty_step": 8}})
    assert s.update_difficulty(0) == 8 and s.update_difficulty(50) == 32 and s.update_difficulty(1000) == 64
    s = CurriculumScheduler({"min_difficulty": 1, "max_difficulty": 3, "schedule_type": "fixed_discrete",
                             "schedule_config": {"difficulty": [1, 2, 3], "max_step": [5, 10]}})
    assert [s.get_difficulty(t) for t in (1, 5, 6, 10, 11)] == [1, 1, 2, 2, 3]
    s = CurriculumScheduler({"min_difficulty": 8, "max_difficulty": 1024, "schedule_type": "fixed_root",
                             "schedule_config": {"total_curriculum_step": 1000, "difficulty_step": 8, "root_degree": 2}})
    assert s.get_difficulty(250) == 512  # sqrt(0.25) = 0.5 -> 8 + 0.5 * 1016 = 516 -> floor to multiple of 8


def test_mmap_indexed_dataset_roundtrip(tmp_path):
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_sampling.indexed_dataset import MMapIndexedDataset, MMapIndexedDatasetBuilder
    prefix = str(tmp_path / "ds")
    b = MMapIndexedDatasetBuilder(prefix + ".bin", dtype=np.int32)
    items = [np.arange(5), np.arange(3) + 10, np.array([7])]
    for it in items:
        b.add_item_numpy(it.astype(np.int32))
        b.end_document()
    b.finalize(prefix + ".idx")
    ds = MMapIndexedDataset(prefix)
    assert len(ds) == 3 and list(ds.sizes) == [5, 3, 1]
    for got, want in zip(ds[0:3], items):
        assert np.array_equal(got, want)
    assert np.array_equal(ds.get(0, offset=2, length=2), [2, 3])


class _Toy(torch.utils.data.Dataset):

    def __init__(self, n):
        g = torch.Generator().manual_seed(0)
        self.lens = torch.randint(1, 9, (n, ), generator=g)

    def __len__(self):
        return len(self.lens)

    def __getitem__(self, i):
        return {"index": torch.tensor([i]), "seqlen": self.lens[i]}


def test_analyzer_then_curriculum_sampler(tmp_path):
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_sampling.data_analyzer import DataAnalyzer
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_sampling.data_sampler import DeepSpeedDataSampler
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_sampling.indexed_dataset import MMapIndexedDataset
    ds = _Toy(200)
    save = str(tmp_path / "analysis")
    for w in range(2):
        DataAnalyzer(ds, num_workers=2, worker_id=w, num_threads=2, batch_size=16, metric_names=["seqlen", "total"],
                     metric_functions=[lambda d: d["seqlen"], lambda d: d["seqlen"].sum().reshape(1)],
                     metric_types=["single_value_per_sample", "accumulate_value_over_samples"], metric_dtypes=[np.int64, np.int64],
                     save_path=save).run_map()
    an = DataAnalyzer(ds, num_workers=2, num_threads=2, metric_names=["seqlen", "total"],
                      metric_types=["single_value_per_sample", "accumulate_value_over_samples"], metric_dtypes=[np.int64, np.int64], save_path=save)
    an.run_reduce()
    i2m = MMapIndexedDataset(f"{save}/seqlen/seqlen_index_to_metric")
    i2s = MMapIndexedDataset(f"{save}/seqlen/seqlen_index_to_sample")
    assert sum(len(r) for r in i2s) == 200
    for row in range(len(i2m)):
        assert all(int(ds.lens[j]) == int(i2m[row][0]) for j in i2s[row])
    assert int(MMapIndexedDataset(f"{save}/total/total_metric_value")[0][0]) == int(ds.lens.sum())

    cfg = {DATA_EFFICIENCY_SEED: 1, DATA_SAMPLING: {DATA_SAMPLING_NUM_EPOCHS: 2, CURRICULUM_LEARNING: {
        CURRICULUM_LEARNING_ENABLED: True, CURRICULUM_LEARNING_CLUSTER_PATH: str(tmp_path / "clusters"),
        CURRICULUM_LEARNING_METRICS: {"seqlen": {
            CURRICULUM_LEARNING_SAMPLE_PATH: f"{save}/seqlen/seqlen_index_to_sample", CURRICULUM_LEARNING_METRIC_PATH: f"{save}/seqlen/seqlen_index_to_metric",
            CURRICULUM_LEARNING_DIFFICULTY_TYPE: "value", CURRICULUM_LEARNING_CLUSTERING_TYPE: "schedule_based", "min_difficulty": 2,
            "max_difficulty": 8, "schedule_type": "fixed_discrete", "schedule_config": {"difficulty": [2, 5, 8], "max_step": [3, 6]}}}}}}
    sampler = DeepSpeedDataSampler(cfg, one_epoch_total_samples=200, micro_batch_size=4, data_parallel_rank=0, data_parallel_size=1,
                                   data_parallel_group=None, gradient_accumulation_steps=2, global_rank=0)
    it = iter(sampler)
    seen_max = []
    for step in range(10):
        batch = next(it) + next(it)  # two micro-batches = one global batch
        assert len(batch) == 8
        seen_max.append(max(int(ds.lens[i]) for i in batch))
        limit = 2 if step < 3 else (5 if step < 6 else 8)
        assert seen_max[-1] <= limit, (step, seen_max)
    sd = sampler.state_dict()
    assert sd[CURRICULUM_LEARNING_STEP] == 10 and len(sd[CURRICULUM_LEARNING_DATA_CLUSTER_PATHS]) == 3
    # resume: a fresh sampler loaded from the state (and the cluster files on disk) continues with the same batches
    import copy
    resumed = DeepSpeedDataSampler(cfg, one_epoch_total_samples=200, micro_batch_size=4, data_parallel_rank=0, data_parallel_size=1,
                                   data_parallel_group=None, gradient_accumulation_steps=2, global_rank=0)
    resumed.load_state_dict(copy.deepcopy(sd))
    it2 = iter(resumed)
    for _ in range(6):
        assert next(it) == next(it2)
    assert resumed.state_dict()[CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION] == sampler.state_dict()[CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION]


def test_random_ltd_layer_cpu():
    from deepspeedsyclsupport_b200.ops.random_ltd import gpt_sample_tokens
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_routing.basic_layer import RandomLayerTokenDrop
    from deepspeedsyclsupport_b200.runtime.data_pipeline.data_routing.scheduler import RandomLTDScheduler
    idx, _ = gpt_sample_tokens(6, 16, 3, layers=2)
    assert idx.shape == (2, 3, 6) and bool((idx[..., 1:] > idx[..., :-1]).all())
    cfg = {RANDOM_LTD_TOTAL_LAYER_NUM: 4, RANDOM_LTD_LAYER_NUM: 1, RANDOM_LTD_GLOBAL_BATCH_SIZE: 2, RANDOM_LTD_MICRO_BATCH_SIZE: 2,
           RANDOM_LTD_MODEL_MASK_NAME: None, RANDOM_LTD_MODEL_TYPE: "decoder", RANDOM_LTD_HIDDEN_STATE_ORDER: "batch_seq_dim",
           RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE: {"enabled": False},
           RANDOM_LTD_SCHEDULER: {RANDOM_LTD_MIN_VALUE: 4, RANDOM_LTD_MAX_VALUE: 16, RANDOM_LTD_SCHEDULER_TYPE: "fixed_linear",
                                  RANDOM_LTD_SCHEDULE_CONFIG: {RANDOM_LTD_REQUIRE_STEP: 10, RANDOM_LTD_INCREASE_STEP: 4}}}
    sched = RandomLTDScheduler(cfg)
    sched.update_seq(0)
    assert sched.get_current_seq() == 4 and sched.state[RANDOM_LTD_CONSUMED_LAYER_TOKENS] == 2 * (4 * 1 + 16 * 3)
    inner = torch.nn.Linear(8, 8)
    layer = RandomLayerTokenDrop(inner)
    layer.init_config(cfg, sched, 0)
    layer.train()
    x = torch.randn(2, 16, 8, requires_grad=True)
    y = layer(x)
    kept = sched.state[RANDOM_LTD_SAMPLE_INDEX][0].long()
    for b in range(2):
        mask = torch.zeros(16, dtype=torch.bool)
        mask[kept[b]] = True
        assert torch.allclose(y[b, mask], inner(x[b, mask]), atol=1e-6)  # kept tokens went through the layer
        assert torch.equal(y[b, ~mask], x[b, ~mask])  # dropped tokens bypass it
    y.sum().backward()
    for b in range(2):
        mask = torch.zeros(16, dtype=torch.bool)
        mask[kept[b]] = True
        assert torch.allclose(x.grad[b, ~mask], torch.ones(14 - 2, 8)) if False else torch.allclose(x.grad[b, ~mask], torch.ones_like(x.grad[b, ~mask]))
        assert torch.allclose(x.grad[b, mask], inner.weight.sum(0).expand(4, 8), atol=1e-6)
    sched.update_seq(10)
    assert sched.get_current_seq() == 16
    layer.eval()
    assert torch.allclose(layer(x), inner(x))
