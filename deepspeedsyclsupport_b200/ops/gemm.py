"""bf16 GEMM on tcgen05 tensor cores (csrc/gemm/gemm_bf16.cu) + autograd-aware linear.

`linear(x, w)` is the replacement for F.linear on the training path (reference:
deepspeed/runtime/zero/linear.py:43-112 LinearFunctionForZeroStage3 -> cuBLAS).  Forward,
dgrad and wgrad each map to one layout variant of the same kernel; wgrad can accumulate
directly into an fp32 main-grad buffer with TMA reduce-add.
"""
import os

import torch

from . import native
from . import recompute as _rc
from ..profiling.flops_profiler import profiler as _prof


def _gemm_ref(a, b, a_mn, b_mn, out_dtype, out=None, accumulate=False):
    A = a.t() if a_mn else a
    Bt = b if b_mn else b.t()
    r = (A.float() @ Bt.float())
    if out is None:
        return r.to(out_dtype)
    if accumulate:
        out.add_(r.to(out.dtype))
    else:
        out.copy_(r.to(out.dtype))
    return out


_SKINNY_ON = os.environ.get("DSB200_SKINNY_GEMM", "1") != "0"
_SKINNY_CLUSTER = os.environ.get("DSB200_SKINNY_CLUSTER", "1") != "0"
_FORCE_SPLITS = int(os.environ.get("DSB200_SKINNY_SPLITS", "0"))  # diagnostics: force the number of K slices
native.register_sigs({"dsb200_gemm_skinny_bf16": "pppp iii iii i p"})


def gemm(a, b, *, a_mn=False, b_mn=False, out=None, out_dtype=torch.bfloat16, accumulate=False, max_ctas=0):
    """C = A @ B^T-style product.

    a: [M,K] (a_mn=False) or stored [K,M] (a_mn=True);  b: [N,K] (b_mn=False) or stored [K,N] (b_mn=True).
    Returns / fills C[M,N].  accumulate=True performs C += A*B (needs `out`).
    """
    if not a.is_cuda:
        return _gemm_ref(a, b, a_mn, b_mn, out_dtype if out is None else out.dtype, out, accumulate)
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "bf16 operands required"
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"K mismatch {K} vs {Kb}"
    if out is None:
        assert not accumulate
        align = 16 // torch.empty((), dtype=out_dtype).element_size()  # TMA needs 16-byte row strides
        ld = (N + align - 1) // align * align
        out = torch.empty(M, ld, dtype=out_dtype, device=a.device)[:, :N]  # e.g. a vocabulary of 50257: padded rows, [M, N] view
    assert out.shape == (M, N) and out.stride(1) == 1
    f32 = out.dtype == torch.float32
    if (M <= 128 and (M <= 32 or N <= 2048) and 512 <= N < 148 * 128 and not (a_mn or b_mn or accumulate or f32) and K % 8 == 0 and N % 4 == 0 and out.stride(0) % 4 == 0 and
            a.stride(0) % 8 == 0 and b.stride(0) % 8 == 0 and _SKINNY_ON):
        # decode-sized problem whose 128 x 256 output tiles would not fill the SMs: weight-streaming kernel (swapped roles, split K).
        # Measured window (profiles/gemm_skinny_bench.json): up to 32 rows, or narrow outputs; with 64+ rows and wide outputs, or
        # >= 148 row tiles of the weight, the tile kernel is as fast or faster.
        tiles, kb = -(-N // 128), -(-K // 64)
        # K slices: enough CTAs to fill the SMs, but every slice keeps >= 32 k-blocks once the row tiles alone cover a quarter of
        # the machine (measured, tools/skinny_splits_probe.py: 64 tiles x K = 3584 / 1024 are fastest unsplit, 56 tiles x K = 8192
        # with 2 slices, 10 tiles x K = 8192 with 8 - 14)
        splits = _FORCE_SPLITS or max(1, min((148 if not max_ctas else max_ctas) // tiles, kb // (32 if tiles >= 32 else 8)))
        # (<= 8 slices reduce inside a thread-block cluster, one kernel; more go through the workspace + reduce kernel, which
        # measured faster than capping a 10-tile problem at 8 slices: profiles/gemm_skinny_bench.json)
        ws = torch.empty(splits * M * N, dtype=torch.float32, device=a.device) if splits > 1 else None
        rc = native.lib().dsb200_gemm_skinny_bf16(native.ptr(a), native.ptr(b), native.ptr(out), native.ptr(ws), M, N, K, a.stride(0), b.stride(0),
                                                  out.stride(0), splits, native.stream_ptr())
        native.check(rc)
        if _prof._ACTIVE:
            _prof.record_flops(2 * M * N * K, M * N * K, "dsb200_gemm_skinny_bf16")
        return out
    mode = (2 if accumulate else 1) if f32 else (3 if accumulate else 0)
    rc = native.lib().dsb200_gemm_bf16(native.ptr(a), native.ptr(b), native.ptr(out), M, N, K, a.stride(0),
                                       b.stride(0), out.stride(0), int(a_mn), int(b_mn), mode, int(max_ctas),
                                       native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        _prof.record_flops(2 * M * N * K, M * N * K, "dsb200_gemm_bf16")
    return out


FP8_MAX = 448.0  # largest finite e4m3 value
_ROWQ_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}
native.register_sigs({"dsb200_fp8_rowwise_quantize": "ppp ii l i p", "dsb200_gemm_fp8_block": "ppppp iii iii i p",
                      "dsb200_gemm_mxfp8": "ppppp iii iii i p", "dsb200_mxfp8_quantize": "ppp ii l i p", "dsb200_mxfp8_sf_bytes": "ii"},
                     restype_long=("dsb200_mxfp8_sf_bytes", ))


def quantize_fp8_rowwise(x):
    """-> (e4m3 tensor of x's shape, fp32 scale per row): x ~= q * scale[:, None].  On the GPU: one kernel, one pass
    (csrc/quant/quantizer.cu fp8_rowwise_quant_kernel); the eager formula below is the CPU path and the oracle."""
    if x.is_cuda and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] % 8 == 0 and x.dtype in _ROWQ_DT and x.data_ptr() % 16 == 0:
        rows, cols = x.shape
        q = torch.empty(rows, cols, dtype=torch.float8_e4m3fn, device=x.device)
        scale = torch.empty(rows, dtype=torch.float32, device=x.device)
        rc = native.lib().dsb200_fp8_rowwise_quantize(native.ptr(x), native.ptr(q), native.ptr(scale), rows, cols, x.stride(0), _ROWQ_DT[x.dtype],
                                                      native.stream_ptr())
        native.check(rc)
        return q, scale
    xf = x.float()
    amax = xf.abs().amax(dim=-1).clamp_min(1e-12)
    scale = amax / FP8_MAX
    q = (xf / scale[:, None]).clamp(-FP8_MAX, FP8_MAX).to(torch.float8_e4m3fn)
    return q, scale


def gemm_fp8(a_q, b_q, a_scale=None, b_scale=None, out=None, max_ctas=0):
    """C[M,N] (bf16) = diag(a_scale) (A B^T) diag(b_scale) on the fp8 tensor-core path (csrc/gemm/gemm_fp8.cu).

    a_q: [M,K] float8_e4m3fn (per-token quantised activations), b_q: [N,K] float8_e4m3fn (per-output-feature quantised
    weights), scales fp32 [M] / [N] (None = 1).  K must be a multiple of 16."""
    assert a_q.dtype == torch.float8_e4m3fn and b_q.dtype == torch.float8_e4m3fn
    M, K = a_q.shape
    N = b_q.shape[0]
    assert b_q.shape[1] == K and a_q.stride(1) == 1 and b_q.stride(1) == 1
    if not a_q.is_cuda:
        r = a_q.float() @ b_q.float().t()
        if a_scale is not None:
            r = r * a_scale[:, None]
        if b_scale is not None:
            r = r * b_scale[None, :]
        return r.to(torch.bfloat16)
    if out is None:
        ld = (N + 7) // 8 * 8
        out = torch.empty(M, ld, dtype=torch.bfloat16, device=a_q.device)[:, :N]
    rc = native.lib().dsb200_gemm_fp8(native.ptr(a_q), native.ptr(b_q), native.ptr(out), native.ptr(a_scale), native.ptr(b_scale), M, N, K,
                                      a_q.stride(0), b_q.stride(0), out.stride(0), int(max_ctas), native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        _prof.record_flops(2 * M * N * K, M * N * K, "dsb200_gemm_fp8")
    return out


def gemm_fp8_block(a_q, b_q, a_scale, b_scale, out=None, max_ctas=0):
    """Block-scaled fp8 GEMM (csrc/gemm/gemm_fp8_block.cu): C[M,N] (bf16) = sum over 128-wide K blocks of
    a_scale[m, kb] * b_scale[n, kb] * (A_kb B_kb^T).  a_q [M,K], b_q [N,K]: float8_e4m3fn or uint8 payloads, K % 128 == 0;
    scales fp32 [M, K/128] / [N, K/128] -- the layout `ops.quantizer.quantize_fp8_block` (the ZeRO++ wire format) produces
    for row-major operands, so a quantised weight shard is consumed as it arrived."""
    M, K = a_q.shape
    N = b_q.shape[0]
    assert b_q.shape[1] == K and K % 128 == 0 and a_q.stride(1) == 1 and b_q.stride(1) == 1
    a_scale, b_scale = a_scale.reshape(M, K // 128), b_scale.reshape(N, K // 128)
    if not a_q.is_cuda:
        af = a_q.view(torch.float8_e4m3fn).float().view(M, K // 128, 128) * a_scale[:, :, None]
        bf = b_q.view(torch.float8_e4m3fn).float().view(N, K // 128, 128) * b_scale[:, :, None]
        return (af.view(M, K) @ bf.view(N, K).t()).to(torch.bfloat16)
    if out is None:
        ld = (N + 7) // 8 * 8
        out = torch.empty(M, ld, dtype=torch.bfloat16, device=a_q.device)[:, :N]
    rc = native.lib().dsb200_gemm_fp8_block(native.ptr(a_q), native.ptr(b_q), native.ptr(out), native.ptr(a_scale.contiguous()),
                                            native.ptr(b_scale.contiguous()), M, N, K, a_q.stride(0), b_q.stride(0), out.stride(0), int(max_ctas),
                                            native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        _prof.record_flops(2 * M * N * K, M * N * K, "dsb200_gemm_fp8_block")
    return out


def linear_fp8_block(x, w_q, w_scale, bias=None):
    """x @ W^T with W held as a block-scaled fp8 payload (one fp32 scale per 128 input features of every output row -- e.g. a
    ZeRO++ qwZ shard, or a checkpoint quantised once at load): activations are quantised to the same format in one kernel and
    the tensor cores multiply fp8 x fp8; the weight is never expanded to 16 bits."""
    from .quantizer import quantize_fp8_block
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    xq, xs = quantize_fp8_block(x2)
    y = gemm_fp8_block(xq.view(x2.shape), w_q, xs, w_scale).view(*x.shape[:-1], w_q.shape[0])
    if y.dtype != x.dtype:
        y = y.to(x.dtype)
    return y if bias is None else y + bias


# ---- MXFP8: e4m3 payload + one UE8M0 (power-of-two) scale per 32 elements, multiplied by the hardware block-scaled MMA
def mxfp8_unpack_scales(sf, rows, K):
    """Packed scale tiles (csrc/gemm/gemm_mxfp8.cu layout) -> fp32 scales [rows, K // 32] (tests, CPU oracle)."""
    KB = K // 128
    r = torch.arange(rows, device=sf.device)[:, None]
    k32 = torch.arange(K // 32, device=sf.device)[None, :]
    idx = ((r // 128) * KB + k32 // 4) * 512 + (r % 32) * 16 + (r % 128 // 32) * 4 + k32 % 4
    return torch.exp2(sf.view(-1)[idx].float() - 127.0)


def _mxfp8_quantize_ref(x):
    rows, K = x.shape
    xf = x.float().view(rows, K // 32, 32)
    amax = xf.abs().amax(-1)
    e = torch.ceil(torch.log2((amax / FP8_MAX).clamp_min(2.0 ** -126))).clamp(-126, 126)
    q = (xf * torch.exp2(-e)[:, :, None]).clamp(-FP8_MAX, FP8_MAX).to(torch.float8_e4m3fn).view(rows, K)
    KB = K // 128
    sf = torch.zeros((rows + 255) // 256 * 2 * KB * 512, dtype=torch.uint8)
    r = torch.arange(rows)[:, None]
    k32 = torch.arange(K // 32)[None, :]
    idx = ((r // 128) * KB + k32 // 4) * 512 + (r % 32) * 16 + (r % 128 // 32) * 4 + k32 % 4
    sf[idx.reshape(-1)] = (e + 127).to(torch.uint8).reshape(-1)
    return q.view(torch.uint8), sf


def quantize_mxfp8(x):
    """x [rows, K] (K % 128 == 0) -> (uint8 e4m3 payload [rows, K], packed UE8M0 scale tiles) for `gemm_mxfp8`."""
    assert x.dim() == 2 and x.shape[1] % 128 == 0 and x.stride(1) == 1
    rows, K = x.shape
    if not x.is_cuda:
        return _mxfp8_quantize_ref(x)
    q = torch.empty(rows, K, dtype=torch.uint8, device=x.device)
    sf = torch.zeros(native.lib().dsb200_mxfp8_sf_bytes(rows, K), dtype=torch.uint8, device=x.device)
    rc = native.lib().dsb200_mxfp8_quantize(native.ptr(x), native.ptr(q), native.ptr(sf), rows, K, x.stride(0), _ROWQ_DT[x.dtype], native.stream_ptr())
    native.check(rc)
    return q, sf


def gemm_mxfp8(a_q, b_q, sfa, sfb, out=None, max_ctas=0):
    """C[M,N] (bf16) = dequant(A) dequant(B)^T on tcgen05.mma.kind::mxf8f6f4.block_scale (csrc/gemm/gemm_mxfp8.cu): the UE8M0 scales
    are applied by the tensor core itself, from tensor memory."""
    M, K = a_q.shape
    N = b_q.shape[0]
    assert b_q.shape[1] == K and K % 128 == 0 and a_q.stride(1) == 1 and b_q.stride(1) == 1
    if not a_q.is_cuda:
        af = a_q.view(torch.float8_e4m3fn).float() * mxfp8_unpack_scales(sfa, M, K).repeat_interleave(32, 1)
        bf = b_q.view(torch.float8_e4m3fn).float() * mxfp8_unpack_scales(sfb, N, K).repeat_interleave(32, 1)
        return (af @ bf.t()).to(torch.bfloat16)
    if out is None:
        ld = (N + 7) // 8 * 8
        out = torch.empty(M, ld, dtype=torch.bfloat16, device=a_q.device)[:, :N]
    rc = native.lib().dsb200_gemm_mxfp8(native.ptr(a_q), native.ptr(b_q), native.ptr(out), native.ptr(sfa), native.ptr(sfb), M, N, K, a_q.stride(0),
                                        b_q.stride(0), out.stride(0), int(max_ctas), native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        _prof.record_flops(2 * M * N * K, M * N * K, "dsb200_gemm_mxfp8")
    return out


def linear_mxfp8(x, w_q, w_sf, bias=None):
    """x @ W^T with W held as MXFP8 (payload + packed scales, quantised once at load); activations are quantised by one kernel."""
    x2 = x.reshape(-1, x.shape[-1])
    x2 = x2 if x2.stride(1) == 1 else x2.contiguous()
    xq, xsf = quantize_mxfp8(x2)
    y = gemm_mxfp8(xq, w_q, xsf, w_sf).view(*x.shape[:-1], w_q.shape[0])
    if y.dtype != x.dtype:
        y = y.to(x.dtype)
    return y if bias is None else y + bias


def linear_fp8(x, w_q, w_scale, bias=None):
    """W8A8 linear for inference: activations quantised per token on the fly, weights pre-quantised per output feature."""
    x2 = x.reshape(-1, x.shape[-1])
    xq, xs = quantize_fp8_rowwise(x2)
    y = gemm_fp8(xq, w_q, xs, w_scale).view(*x.shape[:-1], w_q.shape[0])
    if y.dtype != x.dtype:
        y = y.to(x.dtype)
    return y if bias is None else y + bias


def moe_gemm_ref(x, w, cumsum):
    """fp32 reference: per-expert loop over the row ranges in `cumsum` (host synchronisation)."""
    E, N, K = w.shape
    out = torch.zeros(x.shape[0], N, dtype=x.dtype, device=x.device)
    bounds = [0] + [int(v) for v in cumsum.tolist()]
    for e in range(E):
        a, b = bounds[e], min(bounds[e + 1], x.shape[0])
        if b > a:
            out[a:b] = (x[a:b].float() @ w[e].float().t()).to(x.dtype)
    return out


def moe_gemm(x, w, cumsum, out=None):
    """Grouped expert GEMM (csrc/gemm/gemm_grouped.cu): out[r] = x[r] @ w[expert(r)]^T for rows sorted by expert.

    x: [R, K] bf16;  w: [E, N, K] bf16 (stacked nn.Linear weights);  cumsum: [E] int64 inclusive cumulative rows per
    expert (device tensor -- the kernel reads it, the host never does).  Rows beyond cumsum[-1] are not written.
    (reference: inference/v2/kernels/cutlass_ops/moe_gemm/moe_gemm.py MoEGEMM.__call__)"""
    if not x.is_cuda:
        return moe_gemm_ref(x, w, cumsum)
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and cumsum.dtype == torch.int64
    assert x.dim() == 2 and w.dim() == 3 and x.stride(1) == 1 and w.is_contiguous() and x.shape[1] == w.shape[2]
    R, K = x.shape
    E, N, _ = w.shape
    assert cumsum.numel() == E and cumsum.is_cuda
    if out is None:
        out = torch.empty(R, N, dtype=torch.bfloat16, device=x.device)
    if R == 0:
        return out
    rc = native.lib().dsb200_moe_gemm_bf16(native.ptr(x), native.ptr(w), native.ptr(out), native.ptr(cumsum), R, N, K, E,
                                           x.stride(0), K, out.stride(0), native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        _prof.record_flops(2 * R * N * K, R * N * K, "dsb200_moe_gemm_bf16")
    return out


class _LinearFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, w, w_param, tag, unused_in_recompute):
        ctx.save_for_backward(x, w)
        ctx.w_param = w_param
        x2 = x.reshape(-1, x.shape[-1])
        y = _rc.replayed(tag)  # kept across the checkpointed region: no GEMM during recomputation
        if y is None:
            if unused_in_recompute and _rc.output_unused():
                y = torch.empty(x2.shape[0], w.shape[0], dtype=x.dtype, device=x.device)  # nobody reads it
            else:
                y = gemm(x2, w)
                _rc.record(tag, y)
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        x2 = x.reshape(-1, x.shape[-1])
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn=True).view_as(x)
        if ctx.needs_input_grad[1]:
            wp = ctx.w_param
            if wp is not None:
                # fused gradient accumulation: the wgrad epilogue reduces dW straight into the ZeRO gradient buffer
                mg, fresh = wp.ds_take_main_grad()
                gemm(dy2, x2, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)  # first gradient of the step: plain store
                wp.ds_main_grad_done()
            else:
                dw = gemm(dy2, x2, a_mn=True, b_mn=True)
        return dx, dw, None, None, None


def linear(x, w, tag=None, unused_in_recompute=False):
    """y = x @ w.T with w: [out, in] on the tcgen05 GEMM.  When `w` is managed by the ZeRO store, backward accumulates
    the weight gradient directly into the store's contiguous gradient buffer (no .grad tensor, no copy)."""
    if not x.is_cuda:
        return torch.nn.functional.linear(x, w)
    fused = w.requires_grad and hasattr(w, "ds_get_main_grad") and torch.is_grad_enabled()
    return _LinearFn.apply(x, w, w if fused else None, tag, unused_in_recompute)


class _LinearRopeFn(torch.autograd.Function):
    """QKV projection with RoPE applied to the leading `nrot` heads inside the same autograd node: the rotation is done
    in place on the GEMM output (forward) and on a private copy of the incoming gradient (backward)."""

    @staticmethod
    def forward(ctx, x, w, w_param, inv_freq, positions, seq_len, nrot, head_dim, tag):
        from .fused_ops import _rope_apply
        x2 = x.reshape(-1, x.shape[-1])
        y = _rc.replayed(tag)  # the kept tensor is already rotated
        if y is None:
            y = gemm(x2, w)
            _rope_apply(y, inv_freq, positions, seq_len, 0, nrot, head_dim, False)
            _rc.record(tag, y)
        ctx.save_for_backward(x, w, inv_freq, positions)
        ctx.w_param = w_param
        ctx.rope = (seq_len, nrot, head_dim)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .fused_ops import _rope_apply
        x, w, inv_freq, positions = ctx.saved_tensors
        seq_len, nrot, head_dim = ctx.rope
        dy2 = dy.reshape(-1, dy.shape[-1]).clone()
        _rope_apply(dy2, inv_freq, positions, seq_len, 0, nrot, head_dim, True)
        x2 = x.reshape(-1, x.shape[-1])
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn=True).view_as(x)
        if ctx.needs_input_grad[1]:
            wp = ctx.w_param
            if wp is not None:
                mg, fresh = wp.ds_take_main_grad()
                gemm(dy2, x2, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)  # first gradient of the step: plain store
                wp.ds_main_grad_done()
            else:
                dw = gemm(dy2, x2, a_mn=True, b_mn=True)
        return dx, dw, None, None, None, None, None, None, None


def linear_rope(x, w, inv_freq, seq_len, nrot, head_dim, positions=None, tag=None):
    """[tokens, (hq + 2 hkv) * d] = x @ w.T with rotary embedding applied to heads [0, nrot) (q and k)."""
    if not x.is_cuda:
        from .fused_ops import rope_ref
        y = torch.nn.functional.linear(x, w).reshape(-1, w.shape[0])
        pos = positions if positions is not None else torch.arange(y.shape[0], device=y.device) % seq_len
        return rope_ref(y, inv_freq, pos, 0, nrot, head_dim)
    fused = w.requires_grad and hasattr(w, "ds_get_main_grad") and torch.is_grad_enabled()
    return _LinearRopeFn.apply(x, w, w if fused else None, inv_freq, positions, seq_len, nrot, head_dim, tag)
