"""Autograd-aware wrappers over the element-wise / row-wise CUDA kernels (csrc/norm, csrc/elementwise).

Every function has a pure-PyTorch path used (a) on CPU for plumbing tests and (b) as the fp32 numerical oracle in
the GPU tests.  On a CUDA tensor the native kernel is always used.
"""
import torch
import torch.nn.functional as F

from . import native

_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}
ACT = {"identity": 0, "relu": 1, "gelu": 2, "silu": 3}


def _dt(t):
    return _DT[t.dtype]


# ----------------------------------------------------------------------------- norms
def rms_norm_ref(x, weight, eps, residual=None):
    h = x.float() + (residual.float() if residual is not None else 0)
    y = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps) * weight.float()
    return (y.to(x.dtype), h.to(x.dtype)) if residual is not None else y.to(x.dtype)


class _NormFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, bias, residual, eps, layer_norm, w_param, b_param):
        hidden = x.shape[-1]
        x2 = x.reshape(-1, hidden)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        res_out = torch.empty_like(x2) if residual is not None else None
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device) if layer_norm else None
        r2 = residual.reshape(-1, hidden).contiguous() if residual is not None else None
        rc = native.lib().dsb200_norm_fwd(native.ptr(x2), native.ptr(r2), native.ptr(weight), native.ptr(bias), native.ptr(y),
                                          native.ptr(res_out), native.ptr(mean), native.ptr(rstd), rows, hidden,
                                          float(eps), int(layer_norm), _dt(x), native.stream_ptr())
        native.check(rc)
        ctx.save_for_backward(res_out if residual is not None else x2, weight, mean, rstd)
        ctx.layer_norm = layer_norm
        ctx.has_res = residual is not None
        ctx.has_bias = bias is not None
        ctx.shape = x.shape
        ctx.w_param, ctx.b_param = w_param, b_param
        if residual is not None:
            return y.view(x.shape), res_out.view(x.shape)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy, dres=None):
        xin, weight, mean, rstd = ctx.saved_tensors
        hidden = xin.shape[-1]
        rows = xin.shape[0]
        dy2 = dy.reshape(-1, hidden)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres2 = None
        if dres is not None:
            dres2 = dres.reshape(-1, hidden)
            if not dres2.is_contiguous():
                dres2 = dres2.contiguous()
        dx = torch.empty_like(xin)
        lib = native.lib()
        nscratch = lib.dsb200_norm_bwd_scratch_floats(rows, hidden, int(ctx.layer_norm))
        scratch = torch.empty(nscratch, dtype=torch.float32, device=xin.device)
        use_main = ctx.w_param is not None
        if use_main:
            dgamma = ctx.w_param.ds_get_main_grad()
            dbeta = ctx.b_param.ds_get_main_grad() if ctx.b_param is not None else None
        else:
            dgamma = torch.empty_like(weight)
            dbeta = torch.empty_like(weight) if (ctx.layer_norm and ctx.has_bias) else None
        rc = lib.dsb200_norm_bwd(native.ptr(dy2), native.ptr(xin), native.ptr(weight), native.ptr(mean), native.ptr(rstd),
                                 native.ptr(dres2), native.ptr(dx), native.ptr(dgamma), native.ptr(dbeta), native.ptr(scratch),
                                 rows, hidden, int(ctx.layer_norm), _dt(xin), int(dgamma.dtype == torch.float32), int(use_main),
                                 native.stream_ptr())
        native.check(rc, 2)
        if use_main:
            ctx.w_param.ds_main_grad_done()
            if ctx.b_param is not None:
                ctx.b_param.ds_main_grad_done()
        dx = dx.view(ctx.shape)
        gw = None if use_main else dgamma
        gb = None if use_main else dbeta
        return dx, gw, gb, (dx if ctx.has_res else None), None, None, None, None


def _fusable(p):
    return p is not None and p.requires_grad and hasattr(p, "ds_get_main_grad") and torch.is_grad_enabled()


def rms_norm(x, weight, eps=1e-6, residual=None):
    """y = RMSNorm(x [+ residual]) * weight.  With `residual`, returns (y, x + residual) -- the fused pre-norm form
    (reference: pre_rms_norm, inference/v2/kernels/core_ops/cuda_rms_norm/rms_norm.cu)."""
    if not x.is_cuda:
        return rms_norm_ref(x, weight, eps, residual)
    return _NormFn.apply(x, weight, None, residual, eps, False, weight if _fusable(weight) else None, None)


def layer_norm(x, weight, bias=None, eps=1e-5, residual=None):
    if not x.is_cuda:
        h = x + residual if residual is not None else x
        y = F.layer_norm(h.float(), (x.shape[-1],), weight.float(), bias.float() if bias is not None else None, eps).to(x.dtype)
        return (y, h) if residual is not None else y
    fuse = _fusable(weight) and (bias is None or _fusable(bias))
    return _NormFn.apply(x, weight, bias, residual, eps, True, weight if fuse else None, bias if (fuse and bias is not None) else None)


# ----------------------------------------------------------------------------- RoPE
def rope_inv_freq(head_dim, theta=10000.0, device=None):
    return 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32, device=device) / head_dim))


def rope_ref(x, inv_freq, positions, head0, nheads, head_dim, backward=False):
    """x: [tokens, row] ; rotates heads [head0, head0+nheads) (rotate-half pairing)."""
    T = x.shape[0]
    ang = positions.float()[:, None] * inv_freq[None, :]
    cos, sin = ang.cos()[:, None, :], ang.sin()[:, None, :]
    if backward:
        sin = -sin
    seg = x[:, head0 * head_dim:(head0 + nheads) * head_dim].float().reshape(T, nheads, head_dim)
    a, b = seg[..., :head_dim // 2], seg[..., head_dim // 2:]
    out = torch.cat([a * cos - b * sin, b * cos + a * sin], -1).reshape(T, nheads * head_dim)
    y = x.clone()
    y[:, head0 * head_dim:(head0 + nheads) * head_dim] = out.to(x.dtype)
    return y


def _rope_apply(x2, inv_freq, positions, seq_len, head0, nheads, head_dim, backward):
    rc = native.lib().dsb200_rope_inplace(native.ptr(x2), native.ptr(positions), native.ptr(inv_freq), x2.shape[0], x2.stride(0),
                                          head0, nheads, head_dim, seq_len, int(backward), _dt(x2), native.stream_ptr())
    native.check(rc)


class _RopeFn(torch.autograd.Function):
    """Rotation of the q and k heads of a packed [tokens, (Hq + 2 Hkv) * D] projection (out of place; the in-place
    form lives inside ops.gemm.linear_rope)."""

    @staticmethod
    def forward(ctx, x, inv_freq, positions, seq_len, head0, nheads, head_dim):
        ctx.save_for_backward(inv_freq, positions)
        ctx.args = (seq_len, head0, nheads, head_dim)
        y = x.contiguous().clone()
        _rope_apply(y.view(-1, y.shape[-1]), inv_freq, positions, seq_len, head0, nheads, head_dim, False)
        return y

    @staticmethod
    def backward(ctx, dx):
        inv_freq, positions = ctx.saved_tensors
        seq_len, head0, nheads, head_dim = ctx.args
        dx = dx.contiguous().clone()
        dx2 = dx.view(-1, dx.shape[-1])
        _rope_apply(dx2, inv_freq, positions, seq_len, head0, nheads, head_dim, True)
        return dx, None, None, None, None, None, None


def rope_(x, inv_freq, seq_len, head0, nheads, head_dim, positions=None):
    """Rotate heads [head0, head0+nheads) of x ([..., tokens-major rows]) in place; position = row % seq_len unless
    `positions` (int32 [tokens]) is given."""
    if not x.is_cuda:
        x2 = x.reshape(-1, x.shape[-1])
        pos = positions if positions is not None else torch.arange(x2.shape[0]) % seq_len
        return rope_ref(x2, inv_freq, pos, head0, nheads, head_dim).view(x.shape)
    return _RopeFn.apply(x, inv_freq, positions, seq_len, head0, nheads, head_dim)


# ----------------------------------------------------------------------------- gated activation
def gated_act_ref(x, act="silu", interleaved=False):
    F_ = x.shape[-1] // 2
    g, u = (x[..., 0::2], x[..., 1::2]) if interleaved else (x[..., :F_], x[..., F_:])
    fn = {"silu": F.silu, "gelu": lambda t: F.gelu(t, approximate="tanh"), "relu": F.relu, "identity": lambda t: t}[act]
    return (fn(g.float()) * u.float()).to(x.dtype)


class _GatedActFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, act, interleaved):
        F_ = x.shape[-1] // 2
        x2 = x.reshape(-1, 2 * F_)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        out = torch.empty(x2.shape[0], F_, dtype=x.dtype, device=x.device)
        rc = native.lib().dsb200_gated_act_fwd(native.ptr(x2), native.ptr(None), native.ptr(out), int(x2.shape[0]), F_,
                                               ACT[act], int(interleaved), _dt(x), native.stream_ptr())
        native.check(rc)
        ctx.save_for_backward(x2)
        ctx.act, ctx.interleaved, ctx.shape = act, interleaved, x.shape
        return out.view(*x.shape[:-1], F_)

    @staticmethod
    def backward(ctx, dout):
        (x2,) = ctx.saved_tensors
        F_ = x2.shape[-1] // 2
        d2 = dout.reshape(-1, F_)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        din = torch.empty_like(x2)
        rc = native.lib().dsb200_gated_act_bwd(native.ptr(x2), native.ptr(d2), native.ptr(din), int(x2.shape[0]), F_,
                                               ACT[ctx.act], int(ctx.interleaved), _dt(x2), native.stream_ptr())
        native.check(rc)
        return din.view(ctx.shape), None, None


def gated_act(x, act="silu", interleaved=False):
    """act(gate) * up with [gate | up] packed along the last dim (SiGLU for Llama)."""
    if not x.is_cuda:
        return gated_act_ref(x, act, interleaved)
    return _GatedActFn.apply(x, act, interleaved)


def bias_act(x, bias=None, act="identity"):
    if not x.is_cuda:
        v = x.float() + (bias.float() if bias is not None else 0)
        fn = {"silu": F.silu, "gelu": lambda t: F.gelu(t, approximate="tanh"), "relu": F.relu, "identity": lambda t: t}[act]
        return fn(v).to(x.dtype)
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    out = torch.empty_like(x2)
    rc = native.lib().dsb200_bias_act(native.ptr(x2), native.ptr(bias), native.ptr(out), int(x2.shape[0]), x2.shape[1],
                                      ACT[act], _dt(x), native.stream_ptr())
    native.check(rc)
    return out.view(x.shape)


# ----------------------------------------------------------------------------- cross entropy
class _CrossEntropyFn(torch.autograd.Function):
    """Mean token cross-entropy; backward overwrites the logits buffer with dlogits (no second [T,V] tensor)."""

    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        V = logits.shape[-1]
        l2 = logits.view(-1, V)
        lab = labels.reshape(-1).contiguous()
        rows = l2.shape[0]
        loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
        rc = native.lib().dsb200_cross_entropy_fwd(native.ptr(l2), native.ptr(lab), native.ptr(loss), native.ptr(lse),
                                                   int(rows), V, int(l2.stride(0)),
                                                   int(ignore_index), _dt(l2), native.stream_ptr())
        native.check(rc)
        n_valid = (lab != ignore_index).sum().clamp(min=1).float()
        ctx.save_for_backward(l2, lab, lse, n_valid)
        ctx.ignore_index = ignore_index
        ctx.shape = logits.shape
        return loss.sum() / n_valid

    @staticmethod
    def backward(ctx, g):
        l2, lab, lse, n_valid = ctx.saved_tensors
        V = l2.shape[-1]
        gscale = (g.float() / n_valid).reshape(1).contiguous()
        rc = native.lib().dsb200_cross_entropy_bwd(native.ptr(l2), native.ptr(lab), native.ptr(lse), native.ptr(gscale),
                                                   int(l2.shape[0]), V, int(l2.stride(0)),
                                                   int(ctx.ignore_index), _dt(l2), native.stream_ptr())
        native.check(rc)
        return l2.view(ctx.shape), None, None


def cross_entropy(logits, labels, ignore_index=-100):
    if not logits.is_cuda:
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), labels.reshape(-1), ignore_index=ignore_index)
    return _CrossEntropyFn.apply(logits.contiguous(), labels, ignore_index)


# ----------------------------------------------------------------------------- reductions
def sumsq(x, out=None):
    """sum(x^2) as a 1-element fp32 device tensor in ONE pass over x (bf16 / fp16 / fp32), accumulated into `out` when
    given.  Replaces the .float().square().sum() chain (three full-size temporaries per chunk) in the gradient-norm
    computation (reference: multi_tensor_l2norm / get_grad_norm in runtime/utils.py:315-385)."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
    if not x.is_cuda or x.dtype not in _DT or not x.is_contiguous() or x.data_ptr() % 16 != 0:
        out += x.float().square().sum()
        return out
    native.check(native.lib().dsb200_sumsq(native.ptr(x), x.numel(), _dt(x), native.ptr(out), native.stream_ptr()))
    return out
