"""InferenceEngineV2 -- the FastGen ragged (continuous-batching) engine.
Parity: inference/v2/engine_v2.py:30-254: `put(uids, tokens)` runs one ragged forward and returns last-token logits,
`query(uid, max_request_tokens, max_request_blocks)`, `can_schedule(uids, lengths)`, `flush(uid)`, `serialize()`.
The Dynamic SplitFuse policy itself lives outside the engine (MII in the reference); `scheduler.py` here provides it."""
import json
import os
import pickle
from typing import Iterable, Tuple

import torch

from ... import comm as dist
from ...accelerator import get_accelerator
from ...utils.logging import logger
from .config_v2 import RaggedInferenceEngineConfig
from .ragged.manager_configs import KVCacheConfig
from .ragged.ragged_manager import DSStateManager
from .ragged.ragged_wrapper import RaggedBatchWrapper
from .scheduling_utils import SchedulingError, SchedulingResult


class InferenceEngineV2:

    def __init__(self, model_builder, engine_config: RaggedInferenceEngineConfig):
        """model_builder(tp_group, device) -> RaggedTransformer"""
        self._config = engine_config
        self._initialize_tp_group()
        acc = get_accelerator()
        self.device = torch.device(acc.current_device_name())
        self._model = model_builder(self._tp_group, self.device)
        spec = self._model.spec
        block = int(getattr(engine_config.state_manager, "kv_block_size", 0) or 0) or (128 if self.device.type == "cuda" else 16)
        kv_cfg = KVCacheConfig(block_size=block,
                               cache_shape=(spec.num_layers, self._model.hkv, spec.head_dim), cache_dtype="bf16")
        self._state_manager = DSStateManager(engine_config.state_manager, kv_cfg, base_mp_group=self._tp_group)
        self._batch = RaggedBatchWrapper(engine_config.state_manager)
        self._batch.set_block_table_shape(self._state_manager.max_blocks_per_seq)

    def _initialize_tp_group(self):
        tp = self._config.tensor_parallel.tp_size
        if tp > 1:
            if not dist.is_initialized():
                dist.init_distributed()
            assert dist.get_world_size() >= tp
            self._tp_group = dist.new_group(list(range(tp)))
        else:
            self._tp_group = None

    @property
    def free_blocks(self):
        return self._state_manager.free_blocks

    @property
    def n_kv_cache_groups(self):
        return 1

    def model(self):
        return self._model

    def put(self, batch_uids: Iterable[int], batch_tokens: Iterable[torch.Tensor], do_checks: bool = True) -> torch.Tensor:
        """Run one ragged forward over `batch_tokens[i]` new tokens of sequence `batch_uids[i]`; returns logits
        [n_seqs, vocab] of the last token of every sequence."""
        batch_uids, batch_tokens = list(batch_uids), list(batch_tokens)
        if do_checks:
            token_lens = [t.numel() for t in batch_tokens]
            result = self.can_schedule(batch_uids, token_lens)
            if result != SchedulingResult.Success:
                raise SchedulingError(result)
        self._batch.clear()
        descs = []
        for uid, tokens in zip(batch_uids, batch_tokens):
            seq = self._state_manager.get_or_create_sequence(uid)
            need = seq.kv_blocks_needed(tokens.numel())
            if need > 0:
                seq.extend_kv_cache(self._state_manager.allocate_blocks(need))
            seq.pre_forward(tokens.numel())
            self._batch.insert_sequence(seq, tokens.reshape(-1), do_checks=do_checks)
            descs.append(seq)
        self._batch.finalize()
        with torch.no_grad():
            run = getattr(self._model, "forward_graphed", self._model.forward)  # decode steps replay a CUDA graph
            logits = run(self._batch, self._state_manager)
        for seq in descs:
            seq.post_forward()
        return logits

    def query(self, uid: int, max_request_tokens: int, max_request_blocks: int) -> Tuple[int, int]:
        """How many of `max_request_tokens` can be scheduled for `uid` given `max_request_blocks` free blocks.
        Returns (tokens, blocks needed)."""
        seq = self._state_manager.get_sequence(uid)
        bs = self._state_manager.kv_block_size
        seen = seq.seen_tokens if seq is not None else 0
        allocated = seq.cur_allocated_blocks if seq is not None else 0
        if seq is None and self._state_manager.n_tracked_sequences >= self._config.state_manager.max_tracked_sequences:
            return 0, 0
        free_slots = allocated * bs - seen
        max_tokens_by_blocks = free_slots + max_request_blocks * bs
        tokens = min(max_request_tokens, max_tokens_by_blocks, self._config.state_manager.max_context - seen)
        tokens = max(tokens, 0)
        blocks = max(0, -(-(seen + tokens) // bs) - allocated)
        return tokens, blocks

    def can_schedule(self, uids: Iterable[int], lengths: Iterable[int]) -> SchedulingResult:
        uids, lengths = list(uids), list(lengths)
        cfg = self._config.state_manager
        cur_seqs = self._state_manager.n_tracked_sequences
        free_blocks = self._state_manager.free_blocks
        batch_len = 0
        if len(uids) > cfg.max_ragged_sequence_count:
            return SchedulingResult.BatchSequenceLimitExceeded
        for uid, length in zip(uids, lengths):
            seq = self._state_manager.get_sequence(uid)
            if seq is None:
                cur_seqs += 1
            if cur_seqs > cfg.max_tracked_sequences:
                return SchedulingResult.EngineSequenceLimitExceeded
            sched_len, sched_blocks = self.query(uid, length, free_blocks)
            if sched_len != length:
                return SchedulingResult.KVCacheLimitExceeded
            batch_len += length
            free_blocks -= sched_blocks
        if batch_len > cfg.max_ragged_batch_size:
            return SchedulingResult.BatchTokenLimitExceeded
        return SchedulingResult.Success

    def get_remaining_block_capacity(self, uid: int) -> int:
        seq = self._state_manager.get_sequence(uid)
        if seq is None:
            return 0
        return seq.cur_allocated_blocks * self._state_manager.kv_block_size - seq.seen_tokens

    def flush(self, uid: int) -> None:
        self._state_manager.flush_sequence(uid)

    def serialize(self, save_path: str) -> None:
        """Write this TP rank's flattened parameters + their description; rank 0 also writes the model spec
        (parity: engine_v2.serialize :237-254)."""
        from .model_implementations.flat_model_helpers import (flatten_inference_model, make_metadata_filename, make_model_config_filename,
                                                               make_param_filename)
        os.makedirs(save_path, exist_ok=True)
        m = self._model
        buffer, metadata = flatten_inference_model(m.named_containers(), type(m).__name__)
        torch.save(buffer.cpu(), make_param_filename(save_path, m.tp_rank, m.tp_size))
        with open(make_metadata_filename(save_path, m.tp_rank, m.tp_size), "w") as f:
            json.dump(metadata.model_dump(), f)
        if m.tp_rank == 0:
            with open(make_model_config_filename(save_path), "wb") as f:
                pickle.dump({"spec": m.spec, "dtype": str(m.dtype), "quantization_mode": m.quantization_mode}, f)
