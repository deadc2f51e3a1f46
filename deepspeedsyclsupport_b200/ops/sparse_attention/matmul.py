"""Block-sparse matrix products for attention.

A block-sparse tensor is stored compactly as [batch, nnz, block, block] where nnz is the number of non-zero blocks of
the layout, ordered head-major then row-major (torch.nonzero order) -- the same storage convention as the reference's
Triton kernels (deepspeed/ops/sparse_attention/matmul.py), so modules built on MatMul/Softmax interoperate.

B200 implementation: the non-zero blocks are turned into one batched GEMM (cuBLAS bmm on tensor cores) over gathered
operand blocks; 'dsd'/'dds' results are accumulated into the dense output with index_add.  Compute and memory are
O(nnz * block^2 * d), i.e. genuinely sparse.  The index tables ("LUT") are built once per layout and cached per device.
Modes: sdd  = dense x dense -> sparse   (scores  = Q K^T restricted to the layout)
       dsd  = sparse x dense -> dense   (context = P V)
       dds  = dense x sparse -> dense"""
import torch


def build_lut(layout):
    """(head, row block, col block) of every non-zero block + CSR row pointer over the H*M block rows, built by the host
    library in one pass (csrc/host/sparse_lut.cpp; reference: csrc/sparse_attention/utils.cpp sdd_segment)."""
    import ctypes
    from .. import native
    lay = layout.detach().to("cpu", torch.int64).contiguous()
    H, M, N = lay.shape
    fn = native.host_lib().dsbh_sparse_lut
    fn.restype = ctypes.c_longlong
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 4
    row_ptr = torch.empty(H * M + 1, dtype=torch.int64)
    nnz = fn(lay.data_ptr(), H, M, N, None, None, None, row_ptr.data_ptr())
    h, r, c = (torch.empty(nnz, dtype=torch.int64) for _ in range(3))
    fn(lay.data_ptr(), H, M, N, h.data_ptr(), r.data_ptr(), c.data_ptr(), row_ptr.data_ptr())
    return h, r, c, row_ptr


class _Lut(object):

    def __init__(self, layout, block):
        self.layout = layout
        self.block = block
        self.h, self.r, self.c, self.row_ptr = build_lut(layout)  # sorted head / row / col
        self.nnz = self.h.numel()
        self.cache = {}

    def on(self, device):
        if device not in self.cache:
            self.cache[device] = (self.h.to(device), self.r.to(device), self.c.to(device))
        return self.cache[device]


def _blocks(x, block):
    """[B, H, S, D] -> [B, H, S/block, block, D]"""
    B, H, S, D = x.shape
    return x.view(B, H, S // block, block, D)


class MatMul(object):

    def __init__(self, layout, block, mode, trans_a=False, trans_b=False, bench=False):
        if mode not in ("sdd", "dsd", "dds"):
            raise NotImplementedError("Supported modes are: sdd, dsd, dds")
        self.layout = layout
        self.block = block
        self.mode = mode
        self.trans_a = trans_a
        self.trans_b = trans_b
        self.spdims = layout.shape
        self.lut = _Lut(layout, block)
        self.bench = bench
        self.time_c = self.time_da = self.time_db = None

    def __call__(self, a, b):
        h, r, c = self.lut.on(a.device)
        blk = self.block
        if self.mode == "sdd":
            # a: [B,H,S,D] (or transposed), b: [B,H,S,D] with trans_b (K) or [B,H,D,S]
            A = a.transpose(-1, -2) if self.trans_a else a
            Bm = b if self.trans_b else b.transpose(-1, -2)  # -> [B,H,S,D] so that scores = A @ Bm^T
            Ab = _blocks(A, blk)[:, h, r]  # [B, nnz, blk, D]
            Bb = _blocks(Bm.contiguous(), blk)[:, h, c]  # [B, nnz, blk, D]
            return torch.matmul(Ab, Bb.transpose(-1, -2))  # [B, nnz, blk, blk]
        if self.mode == "dsd":
            # a: sparse [B, nnz, blk, blk], b: dense [B,H,S,D]
            S_ = a.transpose(-1, -2) if self.trans_a else a
            rows, cols = (c, r) if self.trans_a else (r, c)
            Bm = b.transpose(-1, -2) if self.trans_b else b
            Bb = _blocks(Bm.contiguous(), blk)[:, h, cols]  # [B, nnz, blk, D]
            part = torch.matmul(S_, Bb)  # [B, nnz, blk, D]
            Bsz, H, S, D = Bm.shape
            out = torch.zeros(Bsz, H * (S // blk), blk, D, dtype=part.dtype, device=part.device)
            out.index_add_(1, h * (S // blk) + rows, part)
            return out.view(Bsz, H, S, D)
        # dds: a dense [B,H,S,D'] x sparse b [B, nnz, blk, blk] -> dense: out[..., cols] += a[..., rows-block] @ S
        A = a.transpose(-1, -2) if self.trans_a else a
        S_ = b.transpose(-1, -2) if self.trans_b else b
        rows, cols = (c, r) if self.trans_b else (r, c)
        Bsz, H, M, K = A.shape
        Ab = A.view(Bsz, H, M, K // blk, blk).permute(0, 1, 3, 2, 4)[:, h, rows]  # [B, nnz, M, blk]
        part = torch.matmul(Ab, S_)  # [B, nnz, M, blk]
        ncol = self.spdims[2]
        out = torch.zeros(Bsz, H * ncol, M, blk, dtype=part.dtype, device=part.device)
        out.index_add_(1, h * ncol + cols, part)
        return out.view(Bsz, H, ncol, M, blk).permute(0, 1, 3, 2, 4).reshape(Bsz, H, M, ncol * blk)
