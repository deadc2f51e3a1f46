"""Curriculum-aware distributed batch sampler.

Every global batch is drawn from "data clusters": the set of samples that became eligible when the curriculum
difficulty last changed (value- or percentile-based, one or several metrics).  Rank 0 owns the clusters (memory-mapped
index files under `data_cluster_path`, so a resumed job sees the same shuffles), draws the batch and broadcasts the
sample ids; every data-parallel rank then takes its micro-batch slice.
Parity: deepspeed/runtime/data_pipeline/data_sampling/data_sampler.py:36-338 (same constructor, state_dict keys and
cluster file naming).

Organisation here: one `_Metric` record per curriculum metric (scheduler, kind of difficulty, index files) and one `_Cluster`
per generated cluster (name, memory-mapped ids, read cursor); the attributes the reference exposes (`data_cluster_paths`,
`data_cluster_current_position`, `current_difficulties`, ...) are views of those records."""
import os

import numpy as np
import torch

from .... import comm as dist
from ....accelerator import get_accelerator
from ....utils.logging import logger
from ..constants import *  # noqa: F401,F403
from ..curriculum_scheduler import CurriculumScheduler
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype


class _Metric:
    """Curriculum state of one metric."""

    def __init__(self, name, cfg, load_indices):
        self.name = name
        self.scheduler = CurriculumScheduler(cfg)
        self.value_based = cfg[CURRICULUM_LEARNING_DIFFICULTY_TYPE] == CURRICULUM_LEARNING_VALUE_BASED
        self.difficulty_type = cfg[CURRICULUM_LEARNING_DIFFICULTY_TYPE]
        self.clustering_type = cfg[CURRICULUM_LEARNING_CLUSTERING_TYPE]
        self.single_cluster = self.clustering_type == CURRICULUM_LEARNING_SINGLE_CLUSTER
        self.max_difficulty = cfg.get(CURRICULUM_LEARNING_MAX_DIFFICULTY)
        self.index_to_sample = self.index_to_metric = None
        if load_indices and not self.single_cluster:
            # row i of index_to_sample: the sample ids whose metric equals index_to_metric[i]; rows sorted by metric value
            self.index_to_sample = MMapIndexedDataset(cfg[CURRICULUM_LEARNING_SAMPLE_PATH], skip_warmup=True)
            if self.value_based:
                self.index_to_metric = MMapIndexedDataset(cfg[CURRICULUM_LEARNING_METRIC_PATH], skip_warmup=True)

    @property
    def easiest(self):
        """The difficulty 'before the first step': nothing is eligible below it."""
        return float("-inf") if self.value_based else 0


class _Cluster:
    """One generated cluster: its file name under the cluster root, the shuffled ids (rank 0 only) and the read cursor."""

    def __init__(self, fname, ids=None, position=0):
        self.fname, self.ids, self.position = fname, ids, position

    def __len__(self):
        return 0 if self.ids is None else len(self.ids[0])


class DeepSpeedDataSampler(object):

    def __init__(self, data_efficiency_config, one_epoch_total_samples, micro_batch_size, data_parallel_rank, data_parallel_size,
                 data_parallel_group, gradient_accumulation_steps, global_rank, drop_last=True):
        sampling = data_efficiency_config[DATA_SAMPLING]
        self.data_efficiency_config = data_efficiency_config
        self.one_epoch_total_samples = one_epoch_total_samples
        self.total_samples = one_epoch_total_samples * sampling[DATA_SAMPLING_NUM_EPOCHS]
        self.index_dtype = find_fit_int_dtype(0, one_epoch_total_samples)
        self.micro_batch_size, self.gradient_accumulation_steps = micro_batch_size, gradient_accumulation_steps
        self.data_parallel_rank, self.data_parallel_group = data_parallel_rank, data_parallel_group
        self.micro_batch_times_data_parallel_size = micro_batch_size * data_parallel_size
        self.global_batch_size = self.micro_batch_times_data_parallel_size * gradient_accumulation_steps
        self.global_rank, self.drop_last = global_rank, drop_last
        assert self.total_samples > 0, "no sample to consume: {}".format(self.total_samples)
        assert micro_batch_size > 0 and data_parallel_size > 0
        assert data_parallel_rank < data_parallel_size, \
            "data_parallel_rank should be smaller than data size: {}, {}".format(data_parallel_rank, data_parallel_size)
        self.np_rng = np.random.default_rng(data_efficiency_config[DATA_EFFICIENCY_SEED])
        self.state = {}
        self.batch = []
        self.consumed_samples = 0
        cl = sampling[CURRICULUM_LEARNING]
        self.curriculum_enabled = cl[CURRICULUM_LEARNING_ENABLED]
        self._metrics, self._clusters = {}, []
        if self.curriculum_enabled:
            self.curriculum_step = 0
            self.current_difficulties = {}
            self.data_1epoch_size = None
            self.cluster_root = cl[CURRICULUM_LEARNING_CLUSTER_PATH]
            if global_rank == 0:
                os.makedirs(self.cluster_root, exist_ok=True)
            for name, mcfg in cl[CURRICULUM_LEARNING_METRICS].items():
                self._metrics[name] = _Metric(name, mcfg, load_indices=global_rank == 0)

    # ---- the reference's attribute names, as views of the records
    curriculum_schedulers = property(lambda self: {n: m.scheduler for n, m in self._metrics.items()})
    difficulty_type = property(lambda self: {n: m.difficulty_type for n, m in self._metrics.items()})
    clustering_type = property(lambda self: {n: m.clustering_type for n, m in self._metrics.items()})
    curriculum_index_to_sample = property(lambda self: {n: m.index_to_sample for n, m in self._metrics.items() if m.index_to_sample is not None})
    curriculum_index_to_metric = property(lambda self: {n: m.index_to_metric for n, m in self._metrics.items() if m.index_to_metric is not None})
    data_cluster_paths = property(lambda self: [c.fname for c in self._clusters])
    data_cluster_current_position = property(lambda self: [c.position for c in self._clusters])
    data_clusters = property(lambda self: [c.ids for c in self._clusters])
    data_cluster_sizes = property(lambda self: [len(c) for c in self._clusters])

    def __len__(self):
        return self.total_samples

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        for name, fn in schedule_func_dict.items():
            if name in self._metrics:
                self._metrics[name].scheduler.set_custom_get_difficulty(fn)

    def get_start_end_idx(self):
        lo = self.data_parallel_rank * self.micro_batch_size
        return lo, lo + self.micro_batch_size

    # ---- which samples does a difficulty interval (lo, hi] make eligible?
    def get_sample_based_on_metric_value(self, metric, value_start, value_end):
        m = self._metrics[metric]
        picked = [np.copy(m.index_to_sample[row]) for row in range(len(m.index_to_sample)) if value_start < m.index_to_metric[row] <= value_end]
        return np.concatenate(picked, axis=None) if picked else None

    def get_sample_based_on_metric_percentile(self, metric, percentile_start, percentile_end):
        m = self._metrics[metric]
        rows = m.index_to_sample
        if self.data_1epoch_size is None:
            self.data_1epoch_size = sum(len(r) for r in rows)
        per_point = self.data_1epoch_size // m.max_difficulty
        first = per_point * percentile_start
        last = self.data_1epoch_size if percentile_end == m.max_difficulty else per_point * percentile_end
        picked, before = [], 0  # `before` = samples in the rows already passed; take positions [first, last) of the row-major order
        for row in range(len(rows)):
            n = len(rows[row])
            if before + n > first:
                picked.append(np.copy(rows[row][max(0, first - before):min(n, last - before)]))
            before += n
            if before >= last:
                break
        return np.concatenate(picked, axis=None) if picked else None

    def _eligible(self, m, lo, hi):
        pick = self.get_sample_based_on_metric_value if m.value_based else self.get_sample_based_on_metric_percentile
        return pick(m.name, lo, hi)

    def _store(self, fname, ids):
        """Write a (shuffled) id array as the cluster file `fname` and map it back."""
        path = f"{self.cluster_root}/{fname}"
        builder = create_mmap_dataset_builder(path, self.index_dtype)
        builder.add_item_numpy(ids)
        close_mmap_dataset_builder(builder, path)
        return MMapIndexedDataset(path, skip_warmup=True)

    def _new_cluster_ids(self, previous_difficulties):
        """Sample ids that became eligible with the current difficulties (rank 0)."""
        everything = lambda: np.arange(0, self.one_epoch_total_samples, dtype=self.index_dtype)
        clustered = [m for m in self._metrics.values() if not m.single_cluster]
        if len(clustered) > 1:
            # several metrics: intersection of "everything up to the current difficulty" per metric, minus earlier clusters
            ids = None
            for m in self._metrics.values():
                part = everything() if m.single_cluster else self._eligible(m, m.easiest, self.current_difficulties[m.name])
                ids = part if ids is None else np.intersect1d(ids, part, assume_unique=True)
            for c in self._clusters:
                ids = np.setdiff1d(ids, c.ids[0], assume_unique=True)
            return ids
        ids = everything() if not self._clusters else None
        for m in clustered:  # at most one: the samples between the previous and the current difficulty
            ids = self._eligible(m, previous_difficulties[m.name], self.current_difficulties[m.name])
        return ids

    def get_new_cluster(self, previous_difficulties):
        fname = CURRICULUM_LEARNING_CLUSTER_PREFIX + "".join(f"_{n}{self.current_difficulties[n]}" for n in self._metrics)
        mapped = None
        if self.global_rank == 0:
            ids = self._new_cluster_ids(previous_difficulties)
            what = f"new data cluster (previous_difficulties {previous_difficulties}, current_difficulties {self.current_difficulties})"
            if ids is not None and len(ids) > 0:
                logger.info(f"{what} with size {len(ids)} generated.")
                ids = np.array(ids)
                self.np_rng.shuffle(ids)
                mapped = self._store(fname, ids)
            else:
                logger.info(f"{what} has no matched data thus skipped.")
        if dist.is_initialized():
            dist.barrier(group=self.data_parallel_group)
        if os.path.isfile(f"{self.cluster_root}/{fname}.bin"):  # every rank tracks the cluster; only rank 0 holds its ids
            self._clusters.append(_Cluster(fname, mapped))

    def sample_from_clusters(self):
        """How many samples of the next global batch come from each cluster (proportional to the cluster sizes)."""
        sizes = np.asarray(self.data_cluster_sizes, dtype=np.float64)
        draws = self.np_rng.choice(len(sizes), self.global_batch_size, replace=True, p=list(sizes / sizes.sum()))
        return np.bincount(draws, minlength=len(sizes))

    def reshuffle_clusters(self, cidx):
        c = self._clusters[cidx]
        ids = np.copy(c.ids[0])
        self.np_rng.shuffle(ids)
        c.ids = self._store(c.fname, ids)

    def get_sample_from_cluster(self, cidx, num_samples):
        c = self._clusters[cidx]
        taken = list(np.copy(c.ids[0][c.position:c.position + num_samples]))
        c.position += num_samples
        short = num_samples - len(taken)
        if short > 0:  # the cluster is exhausted: new shuffle, continue from its start
            logger.info(f"reshuffling cluster {cidx}.")
            self.reshuffle_clusters(cidx)
            taken += list(np.copy(c.ids[0][:short]))
            c.position = short
        return taken

    def get_next_global_batch(self):
        if not self.curriculum_enabled:
            return
        self.curriculum_step += 1
        previous, changed = {}, False
        for name, m in self._metrics.items():
            now = m.scheduler.update_difficulty(self.curriculum_step)
            previous[name] = self.current_difficulties.get(name, m.easiest)
            changed |= name not in self.current_difficulties or now != self.current_difficulties[name]
            self.current_difficulties[name] = now
        if changed:
            self.get_new_cluster(previous)
        device = get_accelerator().current_device_name() if get_accelerator().is_available() else "cpu"
        if self.global_rank == 0:
            ids = []
            for cidx, count in enumerate(self.sample_from_clusters()):
                ids += self.get_sample_from_cluster(cidx, count)
            self.np_rng.shuffle(ids)
            batch = torch.tensor(np.array(ids, dtype=np.int64), device=device, dtype=torch.long).view(-1)
        else:
            batch = torch.empty(self.global_batch_size, device=device, dtype=torch.long)
        if dist.is_initialized():
            dist.broadcast(batch, 0, group=self.data_parallel_group)
        self.batch = batch.tolist()

    def __iter__(self):
        step = self.micro_batch_times_data_parallel_size
        while self.consumed_samples <= self.total_samples:
            if not self.batch:
                self.get_next_global_batch()
            now, self.batch = self.batch[:step], self.batch[step:]
            if len(now) == step or (now and not self.drop_last):
                lo, hi = self.get_start_end_idx()
                yield now[lo:hi]
                self.consumed_samples += len(now)

    def state_dict(self):
        return {
            CURRICULUM_LEARNING_BATCH: self.batch,
            CURRICULUM_LEARNING_CONSUMED_SAMPLES: self.consumed_samples,
            CURRICULUM_LEARNING_STEP: self.curriculum_step,
            CURRICULUM_LEARNING_CURRENT_DIFFICULTIES: self.current_difficulties,
            CURRICULUM_LEARNING_DATA_CLUSTER_PATHS: self.data_cluster_paths,
            CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION: self.data_cluster_current_position,
            CURRICULUM_LEARNING_NP_RNG_STATE: self.np_rng.bit_generator.state,
        }

    def load_state_dict(self, state_dict):
        self.batch = state_dict[CURRICULUM_LEARNING_BATCH]
        self.consumed_samples = state_dict[CURRICULUM_LEARNING_CONSUMED_SAMPLES]
        self.curriculum_step = state_dict[CURRICULUM_LEARNING_STEP]
        self.current_difficulties = state_dict[CURRICULUM_LEARNING_CURRENT_DIFFICULTIES]
        rng_state = state_dict[CURRICULUM_LEARNING_NP_RNG_STATE]
        if isinstance(rng_state, dict):
            self.np_rng.bit_generator.state = rng_state
        for name, m in self._metrics.items():
            if name in self.current_difficulties:
                m.scheduler.set_current_difficulty(self.current_difficulties[name])
        # cluster files are looked up under THIS job's cluster root by base name (a checkpoint may come from another directory)
        names = [p.split("/")[-1] for p in state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_PATHS]]
        cursors = state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION]
        self._clusters = []
        for fname, position in zip(names, cursors):
            ids = MMapIndexedDataset(f"{self.cluster_root}/{fname}", skip_warmup=True) if self.global_rank == 0 else None
            self._clusters.append(_Cluster(fname, ids, position))
