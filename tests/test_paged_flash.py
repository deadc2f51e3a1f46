"""Paged flash attention (csrc/attn/paged_flash.cu, tcgen05): oracle = dense fp32 attention over the gathered KV history."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make_batch(seq_new, seq_seen, hq, hkv, d, bs, seed=0):
    """Ragged batch of len(seq_new) sequences: seq_seen[i] tokens already cached, seq_new[i] new ones.  Blocks are handed out in a
    shuffled order so page gathering is really exercised."""
    g = torch.Generator().manual_seed(seed)
    n_seq = len(seq_new)
    max_len = max(a + b for a, b in zip(seq_new, seq_seen))
    max_blocks = -(-max_len // bs) + 1
    num_blocks = n_seq * max_blocks + 3
    perm = torch.randperm(num_blocks, generator=g)
    block_table = perm[:n_seq * max_blocks].view(n_seq, max_blocks).int()
    kv_cache = torch.randn(num_blocks, 2, bs, hkv, d, generator=g).bfloat16()
    tok_seq, tok_pos = [], []
    for s in range(n_seq):
        for i in range(seq_new[s]):
            tok_seq.append(s)
            tok_pos.append(seq_seen[s] + i)
    T = len(tok_seq)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, generator=g).bfloat16()
    dev = "cuda"
    return (qkv.to(dev), kv_cache.to(dev), torch.tensor(tok_seq, dtype=torch.int32, device=dev),
            torch.tensor(tok_pos, dtype=torch.int32, device=dev), block_table.to(dev))


def _check(seq_new, seq_seen, hq, hkv, d, bs, n_splits=None, seed=0, decode=None, alibi=False, window=0):
    from deepspeedsyclsupport_b200.inference.v2 import kernels as K
    qkv, kv, ts, tp, bt = _make_batch(seq_new, seq_seen, hq, hkv, d, bs, seed)
    scale = d ** -0.5
    slopes = (2.0 ** (-8.0 * torch.arange(1, hq + 1) / hq)).float().cuda() if alibi else None  # BLOOM's slopes for a power-of-two head count
    got = K.paged_flash_attention(qkv, kv, ts, tp, bt, hq, hkv, d, scale, n_splits=n_splits, decode=decode, alibi_slopes=slopes, window=window).float()
    want = K.paged_attention_ref(qkv.cpu(), kv.cpu(), ts.cpu(), tp.cpu(), bt.cpu(), hq, hkv, d, scale,
                                 None if slopes is None else slopes.cpu(), window).float().cuda()
    torch.cuda.synchronize()
    err = (got - want).abs().max().item()
    assert torch.isfinite(got).all() and err < 2e-2, (seq_new, seq_seen, hq, hkv, d, bs, n_splits, decode, err)


def test_mixed_prefill_and_decode_gqa():
    # GQA 8:1 (16 tokens per atom): a long prefill chunk, a chunk continuing a cached prefix, and three decode steps
    _check(seq_new=[300, 70, 1, 1, 1], seq_seen=[0, 517, 129, 5, 1000], hq=16, hkv=2, d=128, bs=128)


def test_small_blocks_and_mha():
    _check(seq_new=[130, 1, 33], seq_seen=[64, 255, 0], hq=4, hkv=4, d=64, bs=64, seed=1)   # group 1: 128 tokens per atom
    _check(seq_new=[20, 1], seq_seen=[200, 300], hq=8, hkv=2, d=128, bs=32, seed=2)            # 4 cache blocks per 128-key page


def test_split_kv_decode():
    # decode over long contexts with the KV range of every atom cut into 4 and 7 slices (some slices empty for short sequences)
    _check(seq_new=[1, 1, 1], seq_seen=[2000, 130, 3], hq=8, hkv=1, d=128, bs=128, n_splits=4, seed=3)
    _check(seq_new=[1, 40], seq_seen=[900, 500], hq=8, hkv=1, d=128, bs=128, n_splits=7, seed=4)


def test_transposed_decode_kernel():
    # keys on the MMA M axis, 16 q rows per atom: pure decode batches (GQA 8:1, 4:1, MHA), with and without KV splits ...
    _check(seq_new=[1] * 6, seq_seen=[0, 1, 127, 128, 700, 2049], hq=8, hkv=1, d=128, bs=128, seed=5, decode=True, n_splits=1)
    _check(seq_new=[1] * 5, seq_seen=[3000, 130, 3, 511, 512], hq=16, hkv=4, d=128, bs=64, seed=6, decode=True, n_splits=3)
    _check(seq_new=[1, 1, 1], seq_seen=[300, 5, 1000], hq=4, hkv=4, d=128, bs=128, seed=7, decode=True)
    # ... and forced onto a mixed batch: prefill chunks are cut into atoms of 16 / group tokens, still exact
    _check(seq_new=[37, 1, 9], seq_seen=[100, 640, 0], hq=8, hkv=1, d=128, bs=128, seed=8, decode=True, n_splits=2)
    _check(seq_new=[40, 1], seq_seen=[250, 17], hq=8, hkv=8, d=128, bs=32, seed=9, decode=True)


def test_alibi_slopes_in_both_kernels():
    # score += slope[head] * key position: prefill chunks + decode in the general kernel, pure decode in the transposed one
    _check(seq_new=[150, 1, 20], seq_seen=[0, 300, 77], hq=8, hkv=8, d=64, bs=128, seed=11, alibi=True, decode=False)
    _check(seq_new=[40, 1], seq_seen=[500, 129], hq=8, hkv=2, d=128, bs=64, seed=12, alibi=True, decode=False, n_splits=3)
    _check(seq_new=[1, 1, 1, 1], seq_seen=[0, 130, 1000, 2047], hq=16, hkv=16, d=128, bs=128, seed=13, alibi=True, decode=True)
    _check(seq_new=[1, 1], seq_seen=[700, 31], hq=8, hkv=1, d=128, bs=128, seed=14, alibi=True, decode=True, n_splits=2)


def test_local_window_in_both_kernels():
    # a token sees its last `window` keys only; windows shorter and longer than a 128-key page, with and without KV splits
    _check(seq_new=[200, 1, 30], seq_seen=[0, 400, 250], hq=8, hkv=8, d=64, bs=128, seed=21, window=256, decode=False)
    _check(seq_new=[50, 1], seq_seen=[300, 700], hq=8, hkv=2, d=128, bs=64, seed=22, window=100, decode=False, n_splits=2)
    _check(seq_new=[1, 1, 1], seq_seen=[10, 255, 1500], hq=16, hkv=16, d=128, bs=128, seed=23, window=256, decode=True)
    _check(seq_new=[1, 1], seq_seen=[900, 129], hq=8, hkv=1, d=128, bs=128, seed=24, window=64, decode=True, n_splits=3, alibi=True)


def test_dispatch_uses_flash_kernel_and_falls_back():
    from deepspeedsyclsupport_b200.inference.v2 import kernels as K
    from deepspeedsyclsupport_b200.ops import native
    qkv, kv, ts, tp, bt = _make_batch([5, 1], [10, 140], 8, 2, 128, 128)
    assert K.paged_flash_supported(qkv, kv, 8, 2, 128)
    a = K.paged_attention(qkv, kv, ts, tp, bt, 8, 2, 128).float()
    b = K.paged_attention_ref(qkv.cpu(), kv.cpu(), ts.cpu(), tp.cpu(), bt.cpu(), 8, 2, 128, 128 ** -0.5).float().cuda()
    assert (a - b).abs().max().item() < 2e-2
    # head size 96 is outside the tensor-core kernel: the CUDA-core kernel answers
    qkv, kv, ts, tp, bt = _make_batch([3, 1], [4, 20], 4, 2, 96, 128)
    assert not K.paged_flash_supported(qkv, kv, 4, 2, 96)
    a = K.paged_attention(qkv, kv, ts, tp, bt, 4, 2, 96).float()
    b = K.paged_attention_ref(qkv.cpu(), kv.cpu(), ts.cpu(), tp.cpu(), bt.cpu(), 4, 2, 96, 96 ** -0.5).float().cuda()
    assert (a - b).abs().max().item() < 2e-2
    assert native.launch_count > 0
