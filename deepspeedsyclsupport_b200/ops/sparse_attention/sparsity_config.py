"""Block-sparse attention layouts: a layout is a 0/1 tensor [heads, seq/block, seq/block] saying which key blocks each
query block attends to.  Patterns: Dense, Fixed (Sparse Transformers), Variable, BigBird, block-sparse Longformer and
Local sliding window.

The layouts are built with vectorised index arithmetic on block coordinates (row index r, column index c), e.g. a local
window of w blocks is `r // w == c // w`, a sliding window is `|r - c| <= w // 2`, causal is `c <= r`; only the random
blocks use a Python loop (one `random.sample` per row, as the pattern requires).  Constructor arguments, validation
messages and the resulting layouts follow the reference (deepspeed/ops/sparse_attention/sparsity_config.py)."""
import random

import torch


def _grid(num_blocks):
    r = torch.arange(num_blocks).view(-1, 1)
    c = torch.arange(num_blocks).view(1, -1)
    return r, c


def _check_attention(attention):
    if attention not in ("unidirectional", "bidirectional"):
        raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')


def _check_global_ranges(starts, ends):
    if ends is None:
        return
    if len(starts) != len(ends):
        raise ValueError(f"Global block start indices length, {len(starts)}, must be same as global block end indices length, {len(ends)}!")
    for s, e in zip(starts, ends):
        if s >= e:
            raise ValueError(f"Global block start index, {s}, must be smaller than global block end index, {e}!")


class SparsityConfig:
    """Shared properties of the block-sparse patterns: number of heads, block size, per-head layouts or one for all."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        self.num_heads = num_heads
        self.block = block
        self.different_layout_per_head = different_layout_per_head
        self.num_layout_heads = num_heads if different_layout_per_head else 1

    def setup_layout(self, seq_len):
        if seq_len % self.block != 0:
            raise ValueError(f"Sequence Length, {seq_len}, needs to be dividable by Block size {self.block}!")
        num_blocks = seq_len // self.block
        return torch.zeros((self.num_heads, num_blocks, num_blocks), dtype=torch.int64)

    def check_and_propagate_first_head_layout(self, layout):
        if not self.different_layout_per_head:
            layout[1:self.num_heads, :, :] = layout[0, :, :]
        return layout

    def _finish(self, layout, per_head_fn):
        for h in range(self.num_layout_heads):
            layout[h] = per_head_fn(h, layout.shape[1]).to(layout.dtype)
        return self.check_and_propagate_first_head_layout(layout)


class DenseSparsityConfig(SparsityConfig):
    """All blocks attend to all blocks (for comparison / debugging)."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        super().__init__(num_heads, block, different_layout_per_head)

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        layout[:, :, :] = 1
        return layout


class FixedSparsityConfig(SparsityConfig):
    """`Fixed` pattern of "Generative Modeling with Sparse Transformers" (arXiv:1904.10509): local windows of
    `num_local_blocks` plus, per window, `num_global_blocks` representative blocks everybody (later) attends to."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1, attention="bidirectional",
                 horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_local_blocks = num_local_blocks
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError(f"Number of blocks in a local window, {num_local_blocks}, must be dividable by number of global blocks, {num_global_blocks}!")
        self.num_global_blocks = num_global_blocks
        _check_attention(attention)
        self.attention = attention
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        self.horizontal_global_attention = horizontal_global_attention
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout for all heads! Set different_layout_per_head to True.")
        if num_different_global_patterns > (num_local_blocks // num_global_blocks):
            raise ValueError(f"Number of layout versions (num_different_global_patterns), {num_different_global_patterns}, cannot be larger than number of "
                             f"local window blocks divided by number of global blocks, {num_local_blocks} / {num_global_blocks} = "
                             f"{num_local_blocks // num_global_blocks}!")
        self.num_different_global_patterns = num_different_global_patterns

    def _local(self, n):
        r, c = _grid(n)
        same_window = (r // self.num_local_blocks) == (c // self.num_local_blocks)
        return same_window & (c <= r) if self.attention == "unidirectional" else same_window

    def _global_columns(self, h, n):
        """Start indices of the global column groups for head h: the (1 + h % patterns)-th group from the END of every
        complete window, plus one clipped group for a trailing partial window."""
        w, g = self.num_local_blocks, self.num_global_blocks
        first = w - (1 + h % self.num_different_global_patterns) * g
        end = n - n % w
        starts = list(range(first, end, w))
        if end < n:
            starts.append(min(end + first, n - g))
        return starts

    def _global(self, h, n):
        r, c = _grid(n)
        out = torch.zeros(n, n, dtype=torch.bool)
        for s in self._global_columns(h, n):
            cols = (c >= s) & (c < s + self.num_global_blocks)
            out |= cols & ((r >= s) if self.attention == "unidirectional" else torch.ones_like(r, dtype=torch.bool))
            if self.horizontal_global_attention:
                out |= (r >= s) & (r < s + self.num_global_blocks)
        return out

    def set_local_layout(self, h, layout):
        layout[h] |= self._local(layout.shape[1]).to(layout.dtype)
        return layout

    def set_global_layout(self, h, layout):
        layout[h] |= self._global(h, layout.shape[1]).to(layout.dtype)
        return layout

    def make_layout(self, seq_len):
        return self._finish(self.setup_layout(seq_len), lambda h, n: self._local(n) | self._global(h, n))


def _random_blocks(n, count, causal):
    out = torch.zeros(n, n, dtype=torch.bool)
    if count == 0:
        return out
    if n < count:
        raise ValueError(f"Number of random blocks, {count}, must be smaller than overall number of blocks in a row, {n}!")
    for row in range(n):
        pool = range(0, row + 1) if causal else range(0, n)
        out[row, random.sample(pool, min(count, len(pool)))] = True
    return out


def _sliding(n, window_blocks, causal_clip=False):
    if n < window_blocks:
        raise ValueError(f"Number of sliding window blocks, {window_blocks}, must be smaller than overall number of blocks in a row, {n}!")
    r, c = _grid(n)
    band = (c >= r - window_blocks // 2) & (c <= r + window_blocks // 2)
    return band & (c <= r) if causal_clip else band


class VariableSparsityConfig(SparsityConfig):
    """`Variable` pattern: random blocks + local windows of varying size (the last size repeats) + explicit global
    blocks / block ranges."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0, local_window_blocks=[4], global_block_indices=[0],
                 global_block_end_indices=None, attention="bidirectional", horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks = num_random_blocks
        self.local_window_blocks = local_window_blocks
        self.global_block_indices = global_block_indices
        _check_global_ranges(global_block_indices, global_block_end_indices)
        self.global_block_end_indices = global_block_end_indices
        _check_attention(attention)
        self.attention = attention
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        self.horizontal_global_attention = horizontal_global_attention

    def _window_id(self, n):
        """Window index of every block: windows take the listed sizes in order, then the last size repeats."""
        ids, start, wid = torch.zeros(n, dtype=torch.int64), 0, 0
        sizes = list(self.local_window_blocks)
        while start < n:
            size = sizes[wid] if wid < len(sizes) else sizes[-1]
            ids[start:start + size] = wid
            start += size
            wid += 1
        return ids

    def _local(self, n):
        ids = self._window_id(n)
        r, c = _grid(n)
        same = ids.view(-1, 1) == ids.view(1, -1)
        return same & (c <= r) if self.attention == "unidirectional" else same

    def _global(self, n):
        r, c = _grid(n)
        out = torch.zeros(n, n, dtype=torch.bool)
        ends = self.global_block_end_indices or [s + 1 for s in self.global_block_indices]
        for s, e in zip(self.global_block_indices, ends):
            if s >= n:
                continue
            e = min(e, n)
            cols = (c >= s) & (c < e)
            out |= cols & ((r >= s) if self.attention == "unidirectional" else torch.ones_like(r, dtype=torch.bool))
            if self.horizontal_global_attention:
                out |= (r >= s) & (r < e)
        return out

    def set_random_layout(self, h, layout):
        layout[h] |= _random_blocks(layout.shape[1], self.num_random_blocks, causal=False).to(layout.dtype)
        return layout

    def set_local_layout(self, h, layout):
        layout[h] |= self._local(layout.shape[1]).to(layout.dtype)
        return layout

    def set_global_layout(self, h, layout):
        layout[h] |= self._global(layout.shape[1]).to(layout.dtype)
        return layout

    def make_layout(self, seq_len):
        return self._finish(self.setup_layout(seq_len),
                            lambda h, n: _random_blocks(n, self.num_random_blocks, causal=False) | self._local(n) | self._global(n))


class BigBirdSparsityConfig(SparsityConfig):
    """BigBird (arXiv:2007.14062), ITC variant: random + sliding window + the first `num_global_blocks` blocks global."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1, num_sliding_window_blocks=3, num_global_blocks=1,
                 attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks = num_random_blocks
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.num_global_blocks = num_global_blocks
        _check_attention(attention)
        self.attention = attention

    def _head(self, h, n):
        if n < self.num_global_blocks:
            raise ValueError(f"Number of global blocks, {self.num_global_blocks}, must be smaller than overall number of blocks in a row, {n}!")
        causal = self.attention == "unidirectional"
        r, c = _grid(n)
        out = _random_blocks(n, self.num_random_blocks, causal) | _sliding(n, self.num_sliding_window_blocks)
        out |= (r < self.num_global_blocks) | (c < self.num_global_blocks)
        return out & (c <= r) if causal else out

    def set_random_layout(self, h, layout):
        layout[h] |= _random_blocks(layout.shape[1], self.num_random_blocks, self.attention == "unidirectional").to(layout.dtype)
        return layout

    def set_sliding_window_layout(self, h, layout):
        layout[h] |= _sliding(layout.shape[1], self.num_sliding_window_blocks).to(layout.dtype)
        return layout

    def set_global_layout_itc(self, h, layout):
        n = layout.shape[1]
        r, c = _grid(n)
        layout[h] |= ((r < self.num_global_blocks) | (c < self.num_global_blocks)).to(layout.dtype)
        if self.attention == "unidirectional":
            layout = torch.tril(layout)
        return layout

    def make_layout(self, seq_len):
        return self._finish(self.setup_layout(seq_len), self._head)


class BSLongformerSparsityConfig(SparsityConfig):
    """Block-sparse Longformer (arXiv:2004.05150): sliding window + symmetric global rows/columns."""

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3, global_block_indices=[0],
                 global_block_end_indices=None, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.global_block_indices = global_block_indices
        self.attention = attention
        _check_global_ranges(global_block_indices, global_block_end_indices)
        self.global_block_end_indices = global_block_end_indices

    def _head(self, h, n):
        r, c = _grid(n)
        out = _sliding(n, self.num_sliding_window_blocks)
        ends = self.global_block_end_indices or [s + 1 for s in self.global_block_indices]
        for s, e in zip(self.global_block_indices, ends):
            if s < n:
                e = min(e, n)
                out |= ((r >= s) & (r < e)) | ((c >= s) & (c < e))
        return out & (c <= r) if self.attention == "unidirectional" else out

    def set_sliding_window_layout(self, h, layout):
        layout[h] |= _sliding(layout.shape[1], self.num_sliding_window_blocks).to(layout.dtype)
        return layout

    def set_global_layout(self, h, layout):
        layout[h] |= self._head(h, layout.shape[1]).to(layout.dtype)
        return layout

    def make_layout(self, seq_len):
        return self._finish(self.setup_layout(seq_len), self._head)


class LocalSlidingWindowSparsityConfig(SparsityConfig):
    """Purely local sliding window (causal by default)."""

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        layout[h] |= _sliding(layout.shape[1], self.num_sliding_window_blocks, causal_clip=self.attention != "bidirectional").to(layout.dtype)
        return layout

    def make_layout(self, seq_len):
        return self._finish(self.setup_layout(seq_len), lambda h, n: _sliding(n, self.num_sliding_window_blocks, causal_clip=self.attention != "bidirectional"))
