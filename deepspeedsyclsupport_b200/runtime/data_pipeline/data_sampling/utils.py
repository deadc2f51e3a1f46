"""Small helpers of the data sampler / analyzer (parity: deepspeed/runtime/data_pipeline/data_sampling/utils.py)."""
import math

import numpy as np

from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDatasetBuilder


def find_fit_int_dtype(min_value, max_value):
    if min_value >= 0:
        for dt in (np.uint8, np.uint16, np.uint32):
            if max_value <= np.iinfo(dt).max:
                return dt
        return np.uint64
    for dt in (np.int8, np.int16, np.int32):
        if max_value <= np.iinfo(dt).max and min_value >= np.iinfo(dt).min:
            return dt
    return np.int64


def split_index(start_idx, end_idx, num_partitions):
    partition_size = math.ceil((end_idx - start_idx) / num_partitions)
    return [(x, min(end_idx, x + partition_size)) for x in range(start_idx, end_idx, partition_size)]


def split_dataset(dataset, num_workers, worker_id, num_threads):
    worker_splits = split_index(0, len(dataset), num_workers)
    thread_splits = split_index(worker_splits[worker_id][0], worker_splits[worker_id][1], num_threads)
    return worker_splits, thread_splits


def create_mmap_dataset_builder(fname, dtype):
    logger.info(f"Creating mmap dataset builder at {fname}.")
    return MMapIndexedDatasetBuilder(f"{fname}.bin", dtype=dtype)


def close_mmap_dataset_builder(builder, fname):
    builder.end_document()
    builder.finalize(f"{fname}.idx")
    logger.info(f"Finalized mmap dataset builder at {fname}.")
