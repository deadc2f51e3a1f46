"""RaggedBatchWrapper: builds the device metadata of one ragged (continuous) batch in pinned host shadows and ships it
with one non-blocking copy per buffer.  Parity: inference/v2/ragged/ragged_wrapper.py:31-267 (+ the host atom builder
of kernels/ragged_ops/atom_builder)."""
import ctypes

import torch

from ....accelerator import get_accelerator
from .blocked_allocator import _lib


def build_atoms(seq_start, seq_ntok, seq_seen, q_block=128, kv_block=128):
    """[n_atoms, 6] int32: (seq_idx, q_start, q_len, kv_blocks, total_extent, global_q_idx)."""
    n = len(seq_start)
    a = torch.tensor(seq_start, dtype=torch.int32)
    b = torch.tensor(seq_ntok, dtype=torch.int32)
    c = torch.tensor(seq_seen, dtype=torch.int32)
    max_atoms = int(sum(-(-int(x) // q_block) for x in seq_ntok)) + 1
    out = torch.zeros(max_atoms, 6, dtype=torch.int32)
    cnt = _lib().dsbh_build_atoms(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, q_block, kv_block, out.data_ptr(), max_atoms)
    assert cnt >= 0
    return out[:cnt]


class RaggedBatchWrapper:

    def __init__(self, config):
        self._config = config
        self.max_tokens = config.max_ragged_batch_size
        self.max_seqs = config.max_ragged_sequence_count
        dev = get_accelerator().current_device_name()
        pin = get_accelerator().device_name() == "cuda"

        def host(shape, dtype):
            t = torch.zeros(shape, dtype=dtype)
            return t.pin_memory() if pin else t

        self._h_ids = host(self.max_tokens, torch.int64)
        self._h_tok_seq = host(self.max_tokens, torch.int32)
        self._h_tok_pos = host(self.max_tokens, torch.int32)
        self._h_last = host(self.max_seqs, torch.int32)
        self.max_blocks_per_seq = -(-config.max_context // 1)  # set by the state manager
        self._d_ids = torch.zeros(self.max_tokens, dtype=torch.int64, device=dev)
        self._d_tok_seq = torch.zeros(self.max_tokens, dtype=torch.int32, device=dev)
        self._d_tok_pos = torch.zeros(self.max_tokens, dtype=torch.int32, device=dev)
        self._d_last = torch.zeros(self.max_seqs, dtype=torch.int32, device=dev)
        self._h_block_table = None
        self._d_block_table = None
        self.clear()

    def set_block_table_shape(self, max_blocks_per_seq):
        pin = get_accelerator().device_name() == "cuda"
        self.max_blocks_per_seq = max_blocks_per_seq
        t = torch.zeros(self.max_seqs, max_blocks_per_seq, dtype=torch.int32)
        self._h_block_table = t.pin_memory() if pin else t
        self._d_block_table = torch.zeros_like(t, device=get_accelerator().current_device_name())

    def clear(self):
        self._n_tokens = 0
        self._n_seqs = 0
        self._seq_ntok, self._seq_seen, self._seq_start = [], [], []

    def insert_sequence(self, seq_descriptor, tokens, do_checks=True):
        n = tokens.numel()
        if do_checks:
            if self._n_tokens + n > self.max_tokens:
                raise RuntimeError(f"Ragged batch is full: {self._n_tokens} + {n} > {self.max_tokens}")
            if self._n_seqs + 1 > self.max_seqs:
                raise RuntimeError(f"Ragged batch has too many sequences (> {self.max_seqs})")
        self._h_ids[self._n_tokens:self._n_tokens + n].copy_(tokens.to(torch.int64))
        blocks = seq_descriptor.kv_block_ids()
        self._h_block_table[self._n_seqs, :blocks.numel()].copy_(blocks)
        self._seq_start.append(self._n_tokens)
        self._seq_ntok.append(n)
        self._seq_seen.append(seq_descriptor.seen_tokens)
        self._n_tokens += n
        self._n_seqs += 1

    def finalize(self):
        a = torch.tensor(self._seq_ntok, dtype=torch.int32)
        b = torch.tensor(self._seq_seen, dtype=torch.int32)
        total = _lib().dsbh_pack_batch(a.data_ptr(), b.data_ptr(), self._n_seqs, self._h_tok_seq.data_ptr(), self._h_tok_pos.data_ptr(),
                                       self._h_last.data_ptr())
        assert total == self._n_tokens
        self._d_ids[:self._n_tokens].copy_(self._h_ids[:self._n_tokens], non_blocking=True)
        self._d_tok_seq[:self._n_tokens].copy_(self._h_tok_seq[:self._n_tokens], non_blocking=True)
        self._d_tok_pos[:self._n_tokens].copy_(self._h_tok_pos[:self._n_tokens], non_blocking=True)
        self._d_last[:self._n_seqs].copy_(self._h_last[:self._n_seqs], non_blocking=True)
        self._d_block_table[:self._n_seqs].copy_(self._h_block_table[:self._n_seqs], non_blocking=True)

    # device views
    def input_ids(self):
        return self._d_ids[:self._n_tokens]

    def tok_seq(self):
        return self._d_tok_seq[:self._n_tokens]

    def tok_pos(self):
        return self._d_tok_pos[:self._n_tokens]

    def last_tokens(self):
        return self._d_last[:self._n_seqs]

    def block_table(self):
        return self._d_block_table

    def atoms(self, q_block=128, kv_block=128):
        return build_atoms(self._seq_start, self._seq_ntok, self._seq_seen, q_block, kv_block)

    @property
    def current_tokens(self):
        return self._n_tokens

    @property
    def current_sequences(self):
        return self._n_seqs

    @property
    def is_pure_prefill(self):
        """every sequence of the batch is new (no KV history) -- the dense flash kernel can be used per sequence"""
        return all(s == 0 for s in self._seq_seen)

    # ---- reference-named accessors (inference/v2/ragged/ragged_wrapper.py: tensor_toks / tokens_to_seq / ...)
    @property
    def tensor_toks(self):
        return self._n_tokens

    def tokens_to_seq(self, on_device=True):
        """For every token the index of its sequence in the batch."""
        return self._d_tok_seq[:self._n_tokens] if on_device else self._h_tok_seq[:self._n_tokens]

    def batch_metadata_buffer(self, on_device=True):
        """[2] int32: (tokens, sequences) in flight."""
        t = torch.tensor([self._n_tokens, self._n_seqs], dtype=torch.int32)
        return t.to(self._d_ids.device) if on_device else t

    def inflight_seq_descriptors(self, on_device=True):
        """[n_seqs, 4] int32: (first token of the sequence in the batch, tokens in flight, tokens seen before, 0)."""
        t = torch.zeros(self._n_seqs, 4, dtype=torch.int32)
        if self._n_seqs:
            t[:, 0] = torch.tensor(self._seq_start, dtype=torch.int32)
            t[:, 1] = torch.tensor(self._seq_ntok, dtype=torch.int32)
            t[:, 2] = torch.tensor(self._seq_seen, dtype=torch.int32)
        return t.to(self._d_ids.device) if on_device else t

    def kv_ptrs(self, on_device=True):
        """The block table of the batch ([n_seqs, max_blocks] int32 block ids): the kernels index the cache by block id
        rather than by raw pointer."""
        return self._d_block_table[:self._n_seqs] if on_device else self._h_block_table[:self._n_seqs]

    def masks(self, on_device=True):
        """Placeholder for mask metadata (none: causal masking is derived from tok_pos in the kernels)."""
        return None
