"""Head bookkeeping of a tensor-parallel attention layer (parity: inference/v2/model_implementations/sharding/attn.py:11-73)."""
from typing import Optional, Tuple


def _check_gqa(n_heads_q: int, n_heads_kv: int, num_shards: int) -> None:
    if n_heads_q % n_heads_kv != 0:
        raise ValueError("Must be an even ratio between query and key/value heads.")
    if n_heads_kv >= num_shards and n_heads_kv % num_shards != 0:
        raise ValueError("If parallelizing groups, must be able to evenly distribute them.")
    if n_heads_kv < num_shards and num_shards % n_heads_kv != 0:
        raise ValueError("If splitting a group across multiple shards, we must be able to distribute the groups evenly.")


def get_local_heads(shard_rank: int, num_shards: int, n_heads_q: int, n_heads_kv: Optional[int] = None) -> Tuple[int, int]:
    """(query heads, key/value heads) held by `shard_rank`.  MHA: heads are dealt out, the first `heads % shards` ranks get one
    more.  GQA with at least one KV head per shard: whole groups.  GQA with fewer KV heads than shards: a group is split over
    `shards / kv_heads` consecutive ranks, each holding the group's KV head and its share of the query heads."""
    if n_heads_q < num_shards:
        raise ValueError("There must be at least as many attention heads as there are shards.")
    if n_heads_kv is None or n_heads_kv == n_heads_q:
        base, extra = divmod(n_heads_q, num_shards)
        n = base + (1 if shard_rank < extra else 0)
        return n, n
    _check_gqa(n_heads_q, n_heads_kv, num_shards)
    group = n_heads_q // n_heads_kv
    if n_heads_kv >= num_shards:
        local_kv = n_heads_kv // num_shards
        return local_kv * group, local_kv
    ranks_per_group = num_shards // n_heads_kv
    base, extra = divmod(group, ranks_per_group)
    return base + (1 if shard_rank % ranks_per_group < extra else 0), 1
