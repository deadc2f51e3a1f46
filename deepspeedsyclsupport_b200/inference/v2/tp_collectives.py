"""Tensor-parallel collectives of the FastGen ragged engine over peer memory (one NVSwitch domain).

The reference model implementations call `dist.all_reduce` after the attention output projection and after the MLP down
projection of every layer, and all-gather the vocab-parallel logits (inference/v2/model_implementations/llama_v2/model.py:
151-213).  Here one `RaggedTPCollectives` per model owns a symmetric buffer and offers:

* `row_parallel(x, w, residual)`  = all-reduce(x @ w^T) (+ residual) as ONE kernel: `parallel.fused.FusedGemmAllReduce`
  (csrc/gemm/gemm_ar.cu: the GEMM epilogue publishes partial tiles, reducer warps of the same kernel pull the peers' tiles and
  add the residual).  For many tokens on many ranks the one-shot pull moves (W-1) x the output per rank, so from
  `two_shot_min_rows` rows on (and W >= 4) the reduce-scatter form of the kernel is used -- every rank reduces 1/W of the rows --
  followed by a multicast-store all-gather of the row blocks through the switch (csrc/comm/symm.cu nvls_allgather_kernel).
  Decode-sized calls (<= 32 rows, the window where the weight-streaming GEMM wins) are a read of the weight shard: they take the weight-streaming GEMM (every SM, split K)
  into the symmetric staging area + an in-place switch reduction instead.
* `all_reduce(y)` for outputs produced by another kernel (fp8 / MoE GEMMs): in place through the switch
  (`multimem.ld_reduce` + `multimem.st`), or the one-shot pull kernel without a multicast mapping.
* `gather_logits(local)`: every rank stores its vocabulary slice straight into all peers' logits buffers (multicast store, or
  peers pull), one device barrier, no NCCL.
Every rank of the group must make the same sequence of calls (the engine's ranks run the same batches in lock step).
"""
import torch

from ... import comm as dist
from ...ops.gemm import gemm
from ...parallel.fused import FusedGemmAllReduce
from ...parallel.symm import SymmetricBuffer


def _r256(n):
    return (n + 255) // 256 * 256


class RaggedTPCollectives:

    def __init__(self, group, hidden, max_tokens, max_seqs, vocab_local, two_shot_min_rows=256):
        self.group, self.hidden = group, hidden
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.max_tokens = max_tokens
        self.two_shot_min_rows = two_shot_min_rows
        fused_bytes = _r256(FusedGemmAllReduce.bytes_needed(max_tokens, hidden))
        self._act_off = fused_bytes                                   # [max_tokens, hidden] bf16: all-reduce / two-shot staging
        act_bytes = _r256(max_tokens * hidden * 2)
        self._logit_off = self._act_off + act_bytes                    # [world][max_seqs][vocab_local] bf16
        self._logit_bytes = _r256(max_seqs * vocab_local * 2)
        # epochs of the barriers and of the fused GEMMs live in device memory: a decode step captured into a CUDA graph replays
        self.symm = SymmetricBuffer(group, self._logit_off + self.world * self._logit_bytes).use_device_epochs()
        self.fused = FusedGemmAllReduce(self.symm, max_tokens, hidden, offset=0, device_epochs=True)
        self.symm.barrier()
        self.nvls = bool(self.symm.mc_ptr)
        self.calls = {"fused_one_shot": 0, "fused_two_shot": 0, "skinny_switch_reduce": 0, "all_reduce": 0, "gather_logits": 0}

    # ---- GEMM fused with its all-reduce
    def usable_for(self, x, w):
        return (x.is_cuda and x.dtype == torch.bfloat16 and torch.is_tensor(w) and w.dtype == torch.bfloat16 and w.dim() == 2 and
                w.shape[0] == self.hidden and x.shape[0] <= self.max_tokens and x.shape[-1] % 8 == 0 and w.shape[1] == x.shape[-1])

    def row_parallel(self, x, w, residual=None):
        """sum over ranks of x @ w^T (+ residual): [rows, hidden] bf16."""
        rows = x.shape[0]
        x = x if x.is_contiguous() else x.contiguous()
        W = self.world
        if W >= 4 and self.nvls and rows >= self.two_shot_min_rows and rows % W == 0 and (rows // W * self.hidden * 2) % 16 == 0:
            self.calls["fused_two_shot"] += 1
            per = rows // W
            lo = self.rank * per
            mine = self.fused(x, w, residual=None if residual is None else residual[lo:lo + per], rows=(lo, lo + per))
            self.symm.nvls_all_gather(self._act_off, mine)   # my reduced row block -> every rank's staging area
            self.symm.barrier()
            out = self.symm.view(self._act_off, (rows, self.hidden), torch.bfloat16).clone()
            self.symm.barrier()                               # staging may be overwritten by the next call
            return out
        if rows <= 32 and self.nvls and (rows * self.hidden) % (8 * W) == 0 and w.shape[1] % 8 == 0:
            # decode-sized: the op is a read of the weight shard.  The fused kernel's 128 x 256 output tiles put hidden/256 CTAs
            # on it; the weight-streaming GEMM (csrc/gemm/gemm_skinny.cu) uses every SM and writes straight into the symmetric
            # staging area, which the switch then reduces in place (multimem.ld_reduce + multicast store).
            self.calls["skinny_switch_reduce"] += 1
            stage = self.symm.view(self._act_off, (rows, self.hidden), torch.bfloat16)
            gemm(x, w, out=stage)
            self.symm.nvls_all_reduce_bf16(self._act_off, rows * self.hidden)  # barrier, reduce, barrier: nobody touches `stage` after it
            return stage + residual if residual is not None else stage.clone()
        self.calls["fused_one_shot"] += 1
        return self.fused(x, w, residual=residual)

    # ---- all-reduce of a tensor some other kernel produced
    def all_reduce(self, y):
        self.calls["all_reduce"] += 1
        n = y.numel()
        ok = y.is_cuda and y.dtype == torch.bfloat16 and y.is_contiguous() and n * 2 <= _r256(self.max_tokens * self.hidden * 2)
        if not ok:
            dist.all_reduce(y, group=self.group)
            return y
        stage = self.symm.view(self._act_off, (n, ), torch.bfloat16)
        stage.copy_(y.view(-1))
        if self.nvls and n % (8 * self.world) == 0:
            self.symm.nvls_all_reduce_bf16(self._act_off, n)   # barrier, switch reduce + multicast store, barrier
            y.view(-1).copy_(stage)
            self.symm.barrier()
        else:
            self.symm.barrier()
            self.symm.all_reduce_bf16(y.view(-1), self._act_off, n)
            self.symm.barrier()
        return y

    # ---- vocab-parallel logits
    def gather_logits(self, local):
        """local: [n_seqs, vocab_local] (this rank's vocabulary range) -> [n_seqs, world * vocab_local] on every rank."""
        self.calls["gather_logits"] += 1
        n_seqs, v = local.shape
        nbytes = n_seqs * v * 2
        if not (local.is_cuda and local.dtype == torch.bfloat16 and nbytes <= self._logit_bytes and nbytes % 16 == 0):
            full = torch.empty(self.world, n_seqs, v, dtype=local.dtype, device=local.device)
            dist.all_gather_into_tensor(full, local.contiguous(), group=self.group)
            return full.permute(1, 0, 2).reshape(n_seqs, -1)
        local = local.contiguous()
        full = torch.empty(self.world, n_seqs, v, dtype=local.dtype, device=local.device)
        if self.nvls:
            # slices of `nbytes` land back to back: rank r's at logit_off + r * nbytes in EVERY rank's buffer
            self.symm.nvls_all_gather(self._logit_off, local)
            self.symm.barrier()
            full.view(-1).copy_(self.symm.view(self._logit_off, (self.world * n_seqs * v, ), torch.bfloat16))
        else:
            self.symm.view(self._logit_off, (n_seqs * v, ), torch.bfloat16).copy_(local.view(-1))
            self.symm.barrier()
            self.symm.all_gather(full, self._logit_off, nbytes)
        self.symm.barrier()  # the logits area may be rewritten by the next forward
        return full.permute(1, 0, 2).reshape(n_seqs, -1)
