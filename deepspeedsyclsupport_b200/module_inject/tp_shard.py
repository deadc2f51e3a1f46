"""Uneven head sharding helpers (parity: deepspeed/module_inject/tp_shard.py)."""
num_kv_heads = None


def set_num_kv_heads(num):
    global num_kv_heads
    num_kv_heads = num


def get_num_kv_heads():
    return num_kv_heads


def get_shard_size(total_size, mp_size, rank=None):
    from .. import comm as dist
    global num_kv_heads
    if num_kv_heads is not None and total_size % num_kv_heads == 0 and num_kv_heads % mp_size != 0:
        if rank is None:
            rank = dist.get_rank()
        my_slices = (num_kv_heads // mp_size) + (1 if rank < (num_kv_heads % mp_size) else 0)
        return total_size * my_slices // num_kv_heads
    if total_size >= 64:
        grain_size = total_size // 64
        return (grain_size // mp_size + (1 if rank is not None and rank < (grain_size % mp_size) else 0)) * 64 if total_size % (64 * mp_size) else total_size // mp_size
    return total_size // mp_size + (1 if rank is not None and rank < (total_size % mp_size) else 0)


def get_shard_size_list(total_size, mp_size):
    return [get_shard_size(total_size, mp_size, r) for r in range(mp_size)]
