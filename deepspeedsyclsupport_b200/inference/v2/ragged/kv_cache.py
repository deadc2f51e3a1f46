"""Blocked (paged) KV cache.  Parity: inference/v2/ragged/kv_cache.py:40-213: size = free memory minus a reserve,
agreed across the tensor-parallel group with an all-reduce MIN; layout [layers][blocks][2][block_size][kv_heads][head]."""
import torch

from .... import comm as dist
from ....accelerator import get_accelerator
from ....utils.logging import logger
from .blocked_allocator import BlockedAllocator
from .manager_configs import AllocationMode, KVCacheConfig, MemoryConfig


class BlockedKVCache:

    def __init__(self, config: KVCacheConfig, memory_config: MemoryConfig, mp_group=None, offload=False):
        self._config = config
        num_layers, num_heads, head_size = config.cache_shape
        dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}[config.cache_dtype]
        per_block = 2 * config.block_size * num_heads * head_size * num_layers * torch.empty(0, dtype=dtype).element_size()
        if memory_config.mode in (AllocationMode.RESERVE, AllocationMode.RESERVE.value):
            acc = get_accelerator()
            if acc.device_name() == "cuda":
                acc.empty_cache()
                available = acc.available_memory()
            else:
                available = 2 << 30
            total = max(available - memory_config.size, per_block)
            num_blocks = max(1, int(total // per_block))
            if mp_group is not None and dist.is_initialized() and dist.get_world_size(mp_group) > 1:
                t = torch.tensor([num_blocks], dtype=torch.int64, device=acc.current_device_name())
                dist.all_reduce(t, op=dist.ReduceOp.MIN, group=mp_group)
                num_blocks = int(t.item())
        else:
            num_blocks = int(memory_config.size)
        self._num_blocks = num_blocks
        device = get_accelerator().current_device_name()
        logger.info(f"Allocating KV-cache with {num_blocks} blocks ({per_block * num_blocks / 2 ** 30:.2f} GiB)")
        self._cache = torch.zeros(num_layers, num_blocks, 2, config.block_size, num_heads, head_size, dtype=dtype, device=device)
        self._allocator = BlockedAllocator(num_blocks)

    def reserve(self, num_blocks: int) -> torch.Tensor:
        return self._allocator.allocate(num_blocks)

    def free(self, blocks) -> None:
        self._allocator.free(blocks)

    def get_cache(self, layer: int) -> torch.Tensor:
        return self._cache[layer]

    @property
    def free_blocks(self):
        return self._allocator.free_blocks

    @property
    def num_blocks(self):
        return self._num_blocks

    @property
    def block_size(self):
        return self._config.block_size

    @property
    def num_caches(self):
        """Number of cache groups (one: every layer shares the block table; reference supports several KVCacheConfigs)."""
        return 1

    def offload(self, blocks) -> torch.Tensor:
        """Copy the given blocks of every layer to pinned host memory and return the host tensor
        [layers, n, 2, block, heads, head_size]; the blocks themselves stay allocated (reference: kv_cache.offload)."""
        idx = torch.as_tensor(blocks, dtype=torch.long, device=self._cache.device)
        picked = self._cache.index_select(1, idx)
        host = torch.empty(picked.shape, dtype=picked.dtype, device="cpu",
                           pin_memory=self._cache.is_cuda)
        host.copy_(picked, non_blocking=self._cache.is_cuda)
        if self._cache.is_cuda:
            torch.cuda.current_stream().synchronize()
        return host

    def restore(self, blocks, host) -> None:
        """Inverse of `offload`: write host data back into the given blocks."""
        idx = torch.as_tensor(blocks, dtype=torch.long, device=self._cache.device)
        self._cache.index_copy_(1, idx, host.to(self._cache.device, non_blocking=True))
