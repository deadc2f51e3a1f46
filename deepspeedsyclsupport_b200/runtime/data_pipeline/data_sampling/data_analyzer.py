"""Offline dataset analysis for curriculum learning (map / reduce over a torch Dataset).

map   : every (worker, thread) walks its contiguous slice of the dataset in batches, evaluates the metric functions
        and stores, per metric, ONE array for its slice:
            single_value_per_sample       -> <metric>/worker{w}_thread{t}/values.npy   (metric of each sample, in order)
            accumulate_value_over_samples -> <metric>/worker{w}_thread{t}/accumulated.npy
reduce: slices are concatenated in (worker, thread) order (= sample order), then a single stable argsort groups the
        samples by metric value.  Outputs per metric, in the memory-mapped index format the sampler reads:
            <metric>_sample_to_metric                   one row per sample: its metric value
            <metric>_index_to_metric                    one row per distinct value, ascending
            <metric>_index_to_sample                    row i: ids of the samples whose value is index_to_metric[i]
            <metric>_index_to_sample_percentile_merged  the same ids re-bucketed into ~100 rows of consecutive values
            <metric>_metric_value                       (accumulate type) the summed vector
The constructor / run_map / run_reduce surface and the output file names follow the reference
(data_sampling/data_analyzer.py); the per-value CSV spill files of the reference are replaced by the vectorised
argsort grouping above."""
import os
import threading
import time

import numpy as np
import torch
from torch.utils.data import BatchSampler, DataLoader, SequentialSampler, Subset

from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype, split_dataset

SINGLE = "single_value_per_sample"
ACCUMULATE = "accumulate_value_over_samples"


class DataAnalyzer(object):

    def __init__(self, dataset, num_workers=1, worker_id=0, num_threads=1, num_threads_reduce=1, specific_threads=[], batch_size=1,
                 metric_names=[], metric_functions=[], metric_types=[], metric_dtypes=[], save_path="./", collate_fn=None, custom_map_init=None,
                 custom_map_update=None, custom_map_finalize=None, custom_reduce=None):
        self.dataset = dataset
        self.num_workers = num_workers
        self.worker_id = worker_id
        self.num_threads = num_threads
        self.num_threads_reduce = num_threads_reduce
        self.specific_threads = specific_threads
        self.batch_size = batch_size
        self.metric_names = metric_names
        self.metric_functions = metric_functions
        self.metric_types = metric_types
        self.metric_dtypes = metric_dtypes
        self.save_path = save_path
        self.collate_fn = collate_fn
        self.custom_map_init = custom_map_init
        self.custom_map_update = custom_map_update
        self.custom_map_finalize = custom_map_finalize
        self.custom_reduce = custom_reduce
        for t, d in zip(metric_types, metric_dtypes):
            assert t in (SINGLE, ACCUMULATE), f"unknown metric type {t}"
            assert d not in (np.float64, np.double, np.float32), \
                "Currently floating point metric values are not supported. Please change your metric into integer values (and potentially multiply a larger coefficient to keep the precision)."

    # ------------------------------------------------------------------ paths
    def _slice_dir(self, metric, worker, thread):
        return os.path.join(self.save_path, metric, f"worker{worker}_thread{thread}")

    def _out(self, metric, suffix):
        return os.path.join(self.save_path, metric, f"{metric}_{suffix}")

    # ------------------------------------------------------------------ map
    def init_metric_results(self, thread_id, metric_names, metric_types, metric_dtypes, save_path, worker_id):
        results = []
        for name, kind in zip(metric_names, metric_types):
            d = self._slice_dir(name, worker_id, thread_id)
            os.makedirs(d, exist_ok=True)
            results.append({"dir": d, "kind": kind, "chunks": [], "acc": None})
        return results

    def update_metric_results(self, data, metric_types, metric_functions, metric_results):
        for kind, fn, res in zip(metric_types, metric_functions, metric_results):
            values = fn(data)
            if kind == SINGLE:
                n = data["index"].size()[0] if isinstance(data, dict) and "index" in data else values.numel()
                assert values.numel() == n, "metric_function must return a 1-dim tensor that contains metric values for all samples in the batch"
                res["chunks"].append(values.reshape(-1).to("cpu", torch.int64).numpy())
            else:
                res["acc"] = values.clone() if res["acc"] is None else res["acc"].add_(values)

    def finalize_metric_results(self, metric_types, metric_dtypes, metric_results):
        for kind, res in zip(metric_types, metric_results):
            if kind == SINGLE:
                vals = np.concatenate(res["chunks"]) if res["chunks"] else np.zeros(0, dtype=np.int64)
                np.save(os.path.join(res["dir"], "values.npy"), vals)
            elif res["acc"] is not None:
                np.save(os.path.join(res["dir"], "accumulated.npy"), res["acc"].reshape(-1).to("cpu", torch.int64).numpy())

    def run_map_helper(self, thread_id):
        lo, hi = self.thread_splits[thread_id]
        logger.info(f"worker {self.worker_id} thread {thread_id}: start working on data subset {lo} to {hi}")
        subset = Subset(self.dataset, list(range(lo, hi)))
        sampler = BatchSampler(SequentialSampler(subset), batch_size=self.batch_size, drop_last=False)
        kw = {"collate_fn": self.collate_fn} if self.collate_fn is not None else {}
        loader = DataLoader(subset, batch_sampler=sampler, num_workers=0, pin_memory=False, **kw)
        init, update, fin = (self.custom_map_init or self.init_metric_results, self.custom_map_update or self.update_metric_results,
                             self.custom_map_finalize or self.finalize_metric_results)
        results = init(thread_id, self.metric_names, self.metric_types, self.metric_dtypes, self.save_path, self.worker_id)
        t0, done, total = time.time(), 0, max(1, len(subset))
        for data in loader:
            update(data, self.metric_types, self.metric_functions, results)
            done += self.batch_size
            if done % (self.batch_size * 1000) == 0:
                hrs = (time.time() - t0) / 3600.0
                logger.info(f"worker {self.worker_id} thread {thread_id}: {done} out of {total} processed in {hrs:.2f} hr, "
                            f"estimated to finish in {hrs * total / done - hrs:.2f} hr")
        fin(self.metric_types, self.metric_dtypes, results)
        logger.info(f"worker {self.worker_id} thread {thread_id}: finished")

    def run_map(self):
        self.worker_splits, self.thread_splits = split_dataset(self.dataset, self.num_workers, self.worker_id, self.num_threads)
        todo = list(self.specific_threads) if len(self.specific_threads) > 0 else list(range(len(self.thread_splits)))
        if len(todo) == 1:
            self.run_map_helper(todo[0])
            return
        ts = [threading.Thread(target=self.run_map_helper, args=(t, )) for t in todo]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    # ------------------------------------------------------------------ reduce
    def _write_rows(self, fname, rows, dtype):
        b = create_mmap_dataset_builder(fname, dtype)
        for r in rows:
            b.add_item_numpy(np.asarray(r))
        close_mmap_dataset_builder(b, fname)

    def get_metric_value_percentiles(self, metric_name, num_sample_per_value, total_num_samples):
        logger.info(f"Checking the value percentiles of metric {metric_name}...")
        seen, pct = 0, 5
        for value in sorted(num_sample_per_value):
            seen += num_sample_per_value[value]
            while pct <= 100 and seen >= total_num_samples * pct / 100.0:
                logger.info(f"Metric {metric_name} {pct}th percentile: {value}")
                pct += 5

    def _gather_slices(self, metric, fname, num_workers, num_threads):
        parts = []
        for w in range(num_workers):
            for t in range(num_threads):
                path = os.path.join(self._slice_dir(metric, w, t), fname)
                if os.path.isfile(path):
                    parts.append(np.load(path))
        return parts

    def merge_map_results(self, dataset, metric_names, metric_types, save_path, num_workers, num_threads, num_threads_reduce):
        total = len(dataset)
        sample_dtype = find_fit_int_dtype(0, total - 1)
        logger.info(f"Total number of data samples: {total}. Will use {sample_dtype} to store the sample indexes.")
        for metric, kind in zip(metric_names, metric_types):
            if kind == SINGLE:
                values = np.concatenate(self._gather_slices(metric, "values.npy", num_workers, num_threads))
                assert len(values) == total, "The number of samples in map result files are not correct. It's possible that some map worker didn't finish successfully."
                value_dtype = find_fit_int_dtype(int(values.min()), int(values.max()))
                logger.info(f"Metric {metric} has values between {values.min()} and {values.max()}. Will use {value_dtype} to store the metric values.")
                self._write_rows(self._out(metric, "sample_to_metric"), values.reshape(-1, 1), value_dtype)
                order = np.argsort(values, kind="stable")  # sample ids grouped by value, ids ascending inside a group
                distinct, starts, counts = np.unique(values[order], return_index=True, return_counts=True)
                groups = [order[s:s + c] for s, c in zip(starts, counts)]
                self._write_rows(self._out(metric, "index_to_metric"), distinct.reshape(-1, 1), value_dtype)
                self._write_rows(self._out(metric, "index_to_sample"), groups, sample_dtype)
                step = max(1, len(groups) // 100)
                merged = [np.concatenate(groups[i:i + step]) for i in range(0, len(groups), step)]
                self._write_rows(self._out(metric, "index_to_sample_percentile_merged"), merged, sample_dtype)
                self.get_metric_value_percentiles(metric, {int(v): int(c) for v, c in zip(distinct, counts)}, total)
            else:
                parts = self._gather_slices(metric, "accumulated.npy", num_workers, num_threads)
                acc = np.sum(np.stack(parts, 0), axis=0)
                self._write_rows(self._out(metric, "metric_value"), [acc], find_fit_int_dtype(int(acc.min()), int(acc.max())))

    def run_reduce(self):
        reduce_fn = self.custom_reduce or self.merge_map_results
        reduce_fn(self.dataset, self.metric_names, self.metric_types, self.save_path, self.num_workers, self.num_threads, self.num_threads_reduce)
