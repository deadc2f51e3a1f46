// Host natives of the FastGen ragged engine: KV-block free-list allocator, attention-atom builder and batch metadata
// packing.  Reference counterparts: inference/v2/ragged/blocked_allocator.py:11-105 (linked-list allocator, Python in
// the reference), kernels/ragged_ops/atom_builder/atom_builder.cpp:10-53, ragged/csrc/ragged_ops.cpp.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

#define DSBH_EXPORT extern "C" __attribute__((visibility("default")))

struct BlockedAllocator {
  std::vector<int32_t> next;  // linked free list: next[i] = following free block, -1 = end
  int32_t head;
  int32_t free_blocks;
  int32_t total;
};

DSBH_EXPORT void* dsbh_allocator_create(int num_blocks) {
  if (num_blocks < 1) return nullptr;
  BlockedAllocator* a = new BlockedAllocator();
  a->next.resize(num_blocks);
  for (int i = 0; i < num_blocks; ++i) a->next[i] = i + 1 < num_blocks ? i + 1 : -1;
  a->head = 0;
  a->free_blocks = num_blocks;
  a->total = num_blocks;
  return a;
}
DSBH_EXPORT void dsbh_allocator_destroy(void* h) { delete static_cast<BlockedAllocator*>(h); }
DSBH_EXPORT int dsbh_allocator_free_blocks(void* h) { return static_cast<BlockedAllocator*>(h)->free_blocks; }
// returns 0 on success, -1 when fewer than n blocks are free (nothing is allocated then)
DSBH_EXPORT int dsbh_allocator_allocate(void* h, int n, int32_t* out) {
  BlockedAllocator* a = static_cast<BlockedAllocator*>(h);
  if (n > a->free_blocks) return -1;
  for (int i = 0; i < n; ++i) {
    out[i] = a->head;
    a->head = a->next[a->head];
    a->next[out[i]] = -2;  // marks "allocated"
  }
  a->free_blocks -= n;
  return 0;
}
// returns 0 on success, -1 on an invalid / double free (allocator unchanged)
DSBH_EXPORT int dsbh_allocator_free(void* h, const int32_t* blocks, int n) {
  BlockedAllocator* a = static_cast<BlockedAllocator*>(h);
  for (int i = 0; i < n; ++i) {
    if (blocks[i] < 0 || blocks[i] >= a->total || a->next[blocks[i]] != -2) return -1;
    for (int j = 0; j < i; ++j)
      if (blocks[j] == blocks[i]) return -1;
  }
  for (int i = 0; i < n; ++i) {
    a->next[blocks[i]] = a->head;
    a->head = blocks[i];
  }
  a->free_blocks += n;
  return 0;
}

// Attention atoms: the unit of work of the paged attention prefill kernel -- (sequence, q block of <= q_block tokens).
// atom = {seq_idx, q_start (token index in the ragged batch), q_len, kv_blocks (KV blocks visible to the LAST token of
// the atom), total_extent (kv positions visible to the last token), global_q_idx (position of q_start in its sequence)}
struct Atom {
  int32_t seq_idx, q_start, q_len, kv_blocks, total_extent, global_q_idx;
};
// seq_meta: per sequence (start token in batch, n new tokens, tokens already in the KV cache). Returns the atom count.
DSBH_EXPORT int dsbh_build_atoms(const int32_t* seq_start, const int32_t* seq_ntok, const int32_t* seq_seen, int n_seqs, int q_block,
                                 int kv_block, int32_t* atoms_out, int max_atoms) {
  int n = 0;
  for (int s = 0; s < n_seqs; ++s) {
    for (int off = 0; off < seq_ntok[s]; off += q_block) {
      if (n >= max_atoms) return -1;
      Atom a;
      a.seq_idx = s;
      a.q_start = seq_start[s] + off;
      a.q_len = std::min(q_block, seq_ntok[s] - off);
      a.global_q_idx = seq_seen[s] + off;
      a.total_extent = a.global_q_idx + a.q_len;
      a.kv_blocks = (a.total_extent + kv_block - 1) / kv_block;
      memcpy(atoms_out + static_cast<size_t>(n) * 6, &a, sizeof(a));
      ++n;
    }
  }
  return n;
}

// Pack the per-token metadata of a ragged batch: tok_seq[t], tok_pos[t], last_tok[s].
DSBH_EXPORT int dsbh_pack_batch(const int32_t* seq_ntok, const int32_t* seq_seen, int n_seqs, int32_t* tok_seq, int32_t* tok_pos,
                                int32_t* last_tok) {
  int t = 0;
  for (int s = 0; s < n_seqs; ++s) {
    for (int i = 0; i < seq_ntok[s]; ++i) {
      tok_seq[t] = s;
      tok_pos[t] = seq_seen[s] + i;
      ++t;
    }
    last_tok[s] = t - 1;
  }
  return t;
}
