"""Softmax over the rows of a block-sparse score matrix ([batch, nnz, block, block], see matmul.py), with the optional
relative position embedding / key padding mask / attention mask of the reference op
(deepspeed/ops/sparse_attention/softmax.py).  Implementation: a segment softmax -- row maxima and sums are reduced per
(head, row-block) segment with scatter_reduce, so nothing dense is materialised."""
import torch


class Softmax(object):

    def __init__(self, layout, block, bench=False):
        self.num_blocks = int(layout.sum().item())
        self.spdims = layout.shape
        self.layout = layout
        self.block = block
        self.bench = bench
        nz = layout.nonzero(as_tuple=False)
        self._h, self._r, self._c = nz[:, 0], nz[:, 1], nz[:, 2]
        self._seg = nz[:, 0] * layout.shape[1] + nz[:, 1]  # segment id of every block = (head, row block)
        self.lut_cache = {}
        self.time_y = None

    def make_lut(self, device):
        if device not in self.lut_cache:
            self.lut_cache[device] = tuple(t.to(device) for t in (self._h, self._r, self._c, self._seg))
        return self.lut_cache[device]

    def __call__(self, x, scale=1., rpe=None, key_padding_mask=None, attn_mask=None, key_padding_mask_mode="add", attn_mask_mode="add"):
        for name, t in (("relative position embedding", rpe), ("Attention mask", attn_mask), ("Key padding mask", key_padding_mask)):
            if t is not None and t.dtype != x.dtype:
                raise ValueError(f"{name} must be {x.dtype}")
        h, r, c, seg = self.make_lut(x.device)
        blk = self.block
        Bsz, nnz = x.shape[0], x.shape[1]
        s = x.float() * scale
        if rpe is not None:  # [H, S, S] or [B, H, S, S]
            R = rpe if rpe.dim() == 4 else rpe.unsqueeze(0)
            S_ = R.shape[-1]
            Rb = R.view(R.shape[0], R.shape[1], S_ // blk, blk, S_ // blk, blk).permute(0, 1, 2, 4, 3, 5)[:, h, r, c]
            s = s + Rb.float()
        if attn_mask is not None:  # [S, S]
            S_ = attn_mask.shape[-1]
            Mb = attn_mask.view(S_ // blk, blk, S_ // blk, blk).permute(0, 2, 1, 3)[r, c].float()  # [nnz, blk, blk]
            s = s + Mb if attn_mask_mode == "add" else s.masked_fill(Mb == 0, float("-inf"))
        if key_padding_mask is not None:  # [B, S]
            Kb = key_padding_mask.view(Bsz, -1, blk)[:, c].float().unsqueeze(2)  # [B, nnz, 1, blk]
            s = s + Kb if key_padding_mask_mode == "add" else s.masked_fill(Kb == 0, float("-inf"))
        nseg = self.spdims[0] * self.spdims[1]
        idx = seg.view(1, nnz, 1).expand(Bsz, nnz, blk)
        row_max = torch.full((Bsz, nseg, blk), float("-inf"), device=x.device)
        row_max.scatter_reduce_(1, idx, s.amax(-1), reduce="amax", include_self=True)
        row_max = torch.where(torch.isfinite(row_max), row_max, torch.zeros_like(row_max))
        e = torch.exp(s - row_max.gather(1, idx).unsqueeze(-1))
        row_sum = torch.zeros(Bsz, nseg, blk, device=x.device).scatter_add_(1, idx, e.sum(-1))
        out = e / row_sum.gather(1, idx).unsqueeze(-1).clamp_min(1e-30)
        return out.to(x.dtype)
