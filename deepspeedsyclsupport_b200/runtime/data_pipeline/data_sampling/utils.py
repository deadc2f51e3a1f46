"""Small helpers of the data sampler / analyzer (parity: deepspeed/runtime/data_pipeline/data_sampling/utils.py)."""
import numpy as np

from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDatasetBuilder

_UNSIGNED = (np.uint8, np.uint16, np.uint32, np.uint64)
_SIGNED = (np.int8, np.int16, np.int32, np.int64)


def find_fit_int_dtype(min_value, max_value):
    """Narrowest integer dtype that holds [min_value, max_value] (unsigned when nothing is negative)."""
    for dt in (_UNSIGNED if min_value >= 0 else _SIGNED):
        info = np.iinfo(dt)
        if info.min <= min_value and max_value <= info.max:
            return dt
    return _UNSIGNED[-1] if min_value >= 0 else _SIGNED[-1]


def split_index(start_idx, end_idx, num_partitions):
    """[start, end) cut into consecutive ranges of ceil(len / num_partitions) items (the last may be shorter)."""
    step = -(-(end_idx - start_idx) // num_partitions)
    return [(lo, min(lo + step, end_idx)) for lo in range(start_idx, end_idx, step)]


def split_dataset(dataset, num_workers, worker_id, num_threads):
    """(ranges of all workers, ranges of this worker's threads)"""
    by_worker = split_index(0, len(dataset), num_workers)
    return by_worker, split_index(*by_worker[worker_id], num_threads)


def create_mmap_dataset_builder(fname, dtype):
    logger.info(f"Creating mmap dataset builder at {fname}.")
    return MMapIndexedDatasetBuilder(f"{fname}.bin", dtype=dtype)


def close_mmap_dataset_builder(builder, fname):
    builder.end_document()
    builder.finalize(f"{fname}.idx")
    logger.info(f"Finalized mmap dataset builder at {fname}.")
