"""Curriculum-aware distributed batch sampler.

Every global batch is drawn from "data clusters": the set of samples that became eligible when the curriculum
difficulty last changed (value- or percentile-based, one or several metrics).  Rank 0 owns the clusters (memory-mapped
index files under `data_cluster_path`, so a resumed job sees the same shuffles), draws the batch and broadcasts the
sample ids; every data-parallel rank then takes its micro-batch slice.
Parity: deepspeed/runtime/data_pipeline/data_sampling/data_sampler.py:36-338 (same constructor, state_dict keys and
cluster file naming)."""
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


class DeepSpeedDataSampler(object):

    def __init__(self, data_efficiency_config, one_epoch_total_samples, micro_batch_size, data_parallel_rank, data_parallel_size,
                 data_parallel_group, gradient_accumulation_steps, global_rank, drop_last=True):
        self.data_efficiency_config = data_efficiency_config
        self.one_epoch_total_samples = one_epoch_total_samples
        self.index_dtype = find_fit_int_dtype(0, one_epoch_total_samples)
        sampling = data_efficiency_config[DATA_SAMPLING]
        self.total_samples = one_epoch_total_samples * sampling[DATA_SAMPLING_NUM_EPOCHS]
        self.micro_batch_size = micro_batch_size
        self.data_parallel_rank = data_parallel_rank
        self.data_parallel_group = data_parallel_group
        self.micro_batch_times_data_parallel_size = micro_batch_size * data_parallel_size
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.global_batch_size = self.micro_batch_times_data_parallel_size * gradient_accumulation_steps
        self.global_rank = global_rank
        self.drop_last = drop_last
        self.np_rng = np.random.default_rng(data_efficiency_config[DATA_EFFICIENCY_SEED])
        self.state = {}
        self.batch = []
        self.consumed_samples = 0
        self.curriculum_enabled = sampling[CURRICULUM_LEARNING][CURRICULUM_LEARNING_ENABLED]
        if self.curriculum_enabled:
            cl = sampling[CURRICULUM_LEARNING]
            self.curriculum_step = 0
            self.current_difficulties = {}
            self.data_cluster_paths = []
            self.data_cluster_current_position = []
            self.curriculum_schedulers = {}
            self.curriculum_index_to_sample = {}
            self.curriculum_index_to_metric = {}
            self.difficulty_type = {}
            self.clustering_type = {}
            self.data_1epoch_size = None
            self.cluster_root = cl[CURRICULUM_LEARNING_CLUSTER_PATH]
            if self.global_rank == 0:
                self.data_clusters = []
                self.data_cluster_sizes = []
                os.makedirs(self.cluster_root, exist_ok=True)
            for metric, mcfg in cl[CURRICULUM_LEARNING_METRICS].items():
                self.curriculum_schedulers[metric] = CurriculumScheduler(mcfg)
                self.difficulty_type[metric] = mcfg[CURRICULUM_LEARNING_DIFFICULTY_TYPE]
                self.clustering_type[metric] = mcfg[CURRICULUM_LEARNING_CLUSTERING_TYPE]
                if self.global_rank == 0 and self.clustering_type[metric] != CURRICULUM_LEARNING_SINGLE_CLUSTER:
                    self.curriculum_index_to_sample[metric] = MMapIndexedDataset(mcfg[CURRICULUM_LEARNING_SAMPLE_PATH], skip_warmup=True)
                    if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED:
                        self.curriculum_index_to_metric[metric] = MMapIndexedDataset(mcfg[CURRICULUM_LEARNING_METRIC_PATH], skip_warmup=True)
        assert self.total_samples > 0, "no sample to consume: {}".format(self.total_samples)
        assert self.micro_batch_size > 0
        assert data_parallel_size > 0
        assert self.data_parallel_rank < data_parallel_size, \
            "data_parallel_rank should be smaller than data size: {}, {}".format(self.data_parallel_rank, data_parallel_size)

    def __len__(self):
        return self.total_samples

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        for metric in self.curriculum_schedulers:
            if metric in schedule_func_dict:
                self.curriculum_schedulers[metric].set_custom_get_difficulty(schedule_func_dict[metric])

    def get_start_end_idx(self):
        start_idx = self.data_parallel_rank * self.micro_batch_size
        return start_idx, start_idx + self.micro_batch_size

    # ---- eligibility queries on the analyzer's index files: row i of index_to_sample holds the sample ids whose metric
    # ---- equals index_to_metric[i]; rows are sorted by metric value
    def get_sample_based_on_metric_value(self, metric, value_start, value_end):
        rows = []
        i2s, i2m = self.curriculum_index_to_sample[metric], self.curriculum_index_to_metric[metric]
        for row in range(len(i2s)):
            v = i2m[row]
            if value_start < v <= value_end:
                rows.append(np.copy(i2s[row]))
        return np.concatenate(rows, axis=None) if rows else None

    def get_sample_based_on_metric_percentile(self, metric, percentile_start, percentile_end):
        i2s = self.curriculum_index_to_sample[metric]
        if self.data_1epoch_size is None:
            self.data_1epoch_size = sum(len(x) for x in i2s)
        max_percentile = self.data_efficiency_config[DATA_SAMPLING][CURRICULUM_LEARNING][CURRICULUM_LEARNING_METRICS][metric][
            CURRICULUM_LEARNING_MAX_DIFFICULTY]
        per = self.data_1epoch_size // max_percentile
        start_count, end_count = per * percentile_start, per * percentile_end
        if percentile_end == max_percentile:
            end_count = self.data_1epoch_size
        rows, seen = [], 0
        for row in range(len(i2s)):
            n = len(i2s[row])
            if seen + n > start_count:
                a = max(0, start_count - seen)
                b = n if seen + n <= end_count else end_count - seen
                rows.append(np.copy(i2s[row][a:b]))
            seen += n
            if seen >= end_count:
                break
        return np.concatenate(rows, axis=None) if rows else None

    def _eligible(self, metric, lo, hi):
        if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED:
            return self.get_sample_based_on_metric_value(metric, lo, hi)
        return self.get_sample_based_on_metric_percentile(metric, lo, hi)

    def _write_cluster(self, cluster_path, ids):
        builder = create_mmap_dataset_builder(cluster_path, self.index_dtype)
        builder.add_item_numpy(ids)
        close_mmap_dataset_builder(builder, cluster_path)
        return MMapIndexedDataset(cluster_path, skip_warmup=True)

    def get_new_cluster(self, previous_difficulties):
        cluster_fname = CURRICULUM_LEARNING_CLUSTER_PREFIX
        for metric in self.curriculum_schedulers:
            cluster_fname = f"{cluster_fname}_{metric}{self.current_difficulties[metric]}"
        cluster_path = f"{self.cluster_root}/{cluster_fname}"
        if self.global_rank == 0:
            clustered = [m for m in self.clustering_type if self.clustering_type[m] != CURRICULUM_LEARNING_SINGLE_CLUSTER]
            new_cluster = None
            if len(clustered) > 1:
                # several metrics: eligible set = intersection of "everything up to the current difficulty" per metric,
                # minus what previous clusters already cover
                for metric in self.curriculum_schedulers:
                    if self.clustering_type[metric] == CURRICULUM_LEARNING_SINGLE_CLUSTER:
                        metric_cluster = np.arange(0, self.one_epoch_total_samples, dtype=self.index_dtype)
                    else:
                        lo = float("-inf") if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED else 0
                        metric_cluster = self._eligible(metric, lo, self.current_difficulties[metric])
                    new_cluster = metric_cluster if new_cluster is None else np.intersect1d(new_cluster, metric_cluster, assume_unique=True)
                for cluster in self.data_clusters:
                    new_cluster = np.setdiff1d(new_cluster, cluster[0], assume_unique=True)
            else:
                if len(self.data_clusters) == 0:
                    new_cluster = np.arange(0, self.one_epoch_total_samples, dtype=self.index_dtype)
                for metric in clustered:
                    new_cluster = self._eligible(metric, previous_difficulties[metric], self.current_difficulties[metric])
            if new_cluster is not None and len(new_cluster) > 0:
                logger.info(f"new data cluster (previous_difficulties {previous_difficulties}, current_difficulties "
                            f"{self.current_difficulties}) with size {len(new_cluster)} generated.")
                new_cluster = np.array(new_cluster)
                self.np_rng.shuffle(new_cluster)
                self.data_clusters.append(self._write_cluster(cluster_path, new_cluster))
                self.data_cluster_sizes.append(len(self.data_clusters[-1][0]))
            else:
                logger.info(f"new data cluster (previous_difficulties {previous_difficulties}, current_difficulties "
                            f"{self.current_difficulties}) has no matched data thus skipped.")
        if dist.is_initialized():
            dist.barrier(group=self.data_parallel_group)
        if os.path.isfile(f"{cluster_path}.bin"):
            self.data_cluster_paths.append(cluster_fname)
            self.data_cluster_current_position.append(0)

    def sample_from_clusters(self):
        num_clusters = len(self.data_clusters)
        weight_sum = sum(self.data_cluster_sizes)
        weights = [x / weight_sum for x in self.data_cluster_sizes]
        samples = self.np_rng.choice(num_clusters, self.global_batch_size, replace=True, p=weights)
        return np.bincount(samples, minlength=num_clusters)

    def reshuffle_clusters(self, cidx):
        cluster_path = f"{self.cluster_root}/{self.data_cluster_paths[cidx]}"
        cluster = np.copy(self.data_clusters[cidx][0])
        self.np_rng.shuffle(cluster)
        self.data_clusters[cidx] = self._write_cluster(cluster_path, cluster)

    def get_sample_from_cluster(self, cidx, num_samples):
        start_idx = self.data_cluster_current_position[cidx]
        samples = list(np.copy(self.data_clusters[cidx][0][start_idx:(start_idx + num_samples)]))
        self.data_cluster_current_position[cidx] += num_samples
        if len(samples) < num_samples:
            num_samples_remained = num_samples - len(samples)
            logger.info(f"reshuffling cluster {cidx}.")
            self.reshuffle_clusters(cidx)
            samples += list(np.copy(self.data_clusters[cidx][0][:num_samples_remained]))
            self.data_cluster_current_position[cidx] = num_samples_remained
        return samples

    def get_next_global_batch(self):
        if not self.curriculum_enabled:
            return
        self.curriculum_step += 1
        new_cluster = False
        previous_difficulties = {}
        for metric, sched in self.curriculum_schedulers.items():
            next_difficulty = sched.update_difficulty(self.curriculum_step)
            if metric not in self.current_difficulties or next_difficulty != self.current_difficulties[metric]:
                new_cluster = True
            if metric in self.current_difficulties:
                previous_difficulties[metric] = self.current_difficulties[metric]
            else:
                previous_difficulties[metric] = float("-inf") if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED else 0
            self.current_difficulties[metric] = next_difficulty
        if new_cluster:
            self.get_new_cluster(previous_difficulties)
        device = get_accelerator().current_device_name() if get_accelerator().is_available() else "cpu"
        if self.global_rank == 0:
            samples_per_cluster = self.sample_from_clusters()
            batch = []
            for cidx in range(len(samples_per_cluster)):
                batch += self.get_sample_from_cluster(cidx, samples_per_cluster[cidx])
            self.np_rng.shuffle(batch)
            batch = torch.tensor(np.array(batch, dtype=np.int64), device=device, dtype=torch.long).view(-1)
        else:
            batch = torch.empty(self.global_batch_size, device=device, dtype=torch.long)
        if dist.is_initialized():
            dist.broadcast(batch, 0, group=self.data_parallel_group)
        self.batch = batch.tolist()

    def __iter__(self):
        while self.consumed_samples <= self.total_samples:
            if len(self.batch) == 0:
                self.get_next_global_batch()
            current_batch = self.batch[:self.micro_batch_times_data_parallel_size]
            self.batch = self.batch[self.micro_batch_times_data_parallel_size:]
            if len(current_batch) == self.micro_batch_times_data_parallel_size or (len(current_batch) > 0 and not self.drop_last):
                start_idx, end_idx = self.get_start_end_idx()
                yield current_batch[start_idx:end_idx]
                self.consumed_samples += len(current_batch)
                current_batch = []

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
        self.data_cluster_paths = [p.split("/")[-1] for p in state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_PATHS]]
        self.data_cluster_current_position = state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION]
        rng_state = state_dict[CURRICULUM_LEARNING_NP_RNG_STATE]
        if isinstance(rng_state, dict):
            self.np_rng.bit_generator.state = rng_state
        for metric, sched in self.curriculum_schedulers.items():
            if metric in self.current_difficulties:
                sched.set_current_difficulty(self.current_difficulties[metric])
        if self.global_rank == 0:
            self.data_clusters, self.data_cluster_sizes = [], []
            for cluster_fname in self.data_cluster_paths:
                self.data_clusters.append(MMapIndexedDataset(f"{self.cluster_root}/{cluster_fname}", skip_warmup=True))
                self.data_cluster_sizes.append(len(self.data_clusters[-1][0]))
