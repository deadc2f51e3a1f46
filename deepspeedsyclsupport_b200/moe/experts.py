"""Local experts container (parity: deepspeed/moe/experts.py:10-35).

The reference calls the local experts one after another: 2 x num_local_experts GEMM launches per MoE layer, each over the few
rows one expert received.  Capacity-based dispatch gives every local expert the same number of rows, so when the experts are
plain two-layer MLPs (checked numerically at construction: `fc2(act(fc1(x)))` must reproduce the expert's own forward) both
projections run as ONE grouped GEMM over all local experts (csrc/gemm/gemm_grouped.cu for bf16 on the GPU, batched matmul
otherwise), forward and backward.  Anything else (custom experts, dropout, ZeRO-3 sharded expert weights) takes the loop."""
import copy
import os

import torch
import torch.nn.functional as F


class _GroupedLinearFn(torch.autograd.Function):
    """y[e] = x[e] @ W[e]^T + b[e] for x [E, R, K], W [E, N, K], b [E, N] or None."""

    @staticmethod
    def _use_kernel(x, W):
        return x.is_cuda and x.dtype == torch.bfloat16 and W.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and W.shape[1] % 8 == 0

    @staticmethod
    def _cumsum(E, R, device):
        return torch.arange(1, E + 1, device=device, dtype=torch.int64) * R

    @staticmethod
    def _grouped(x3, W):
        """x3 [E, R, K] @ W[E, N, K]^T -> [E, R, N]"""
        E, R, K = x3.shape
        if _GroupedLinearFn._use_kernel(x3, W):
            from ..ops.gemm import moe_gemm
            return moe_gemm(x3.reshape(E * R, K), W.contiguous(), _GroupedLinearFn._cumsum(E, R, x3.device)).view(E, R, W.shape[1])
        return torch.bmm(x3, W.transpose(1, 2))

    @staticmethod
    def forward(ctx, x3, W, b):
        ctx.save_for_backward(x3, W)
        ctx.has_bias = b is not None
        y = _GroupedLinearFn._grouped(x3.contiguous(), W)
        return y if b is None else y + b[:, None, :]

    @staticmethod
    def backward(ctx, dy):
        x3, W = ctx.saved_tensors
        dy = dy.contiguous()
        dx = dW = db = None
        if ctx.needs_input_grad[0]:
            dx = _GroupedLinearFn._grouped(dy, W.transpose(1, 2).contiguous())  # dy [E,R,N] @ (W^T)[E,K,N]^T
        if ctx.needs_input_grad[1]:
            dW = torch.bmm(dy.transpose(1, 2), x3)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.sum(1)
        return dx, dW, db


_ACTIVATIONS = (("gelu", F.gelu), ("gelu_tanh", lambda t: F.gelu(t, approximate="tanh")), ("relu", F.relu), ("silu", F.silu),
                ("identity", lambda t: t))


def _probe_two_layer_mlp(expert):
    """(fc1, fc2, activation) if `expert` computes fc2(act(fc1(x))) with two nn.Linear layers and nothing else, else None."""
    linears = [m for m in expert.modules() if isinstance(m, torch.nn.Linear)]
    if len(linears) != 2:
        return None
    own = {id(p) for m in linears for p in m.parameters()}
    if any(id(p) not in own for p in expert.parameters()) or any(p.is_meta or p.numel() == 0 for p in expert.parameters()):
        return None
    if any(isinstance(m, torch.nn.Dropout) and m.p > 0 for m in expert.modules()):
        return None
    fc1, fc2 = linears if linears[0].out_features == linears[1].in_features else linears[::-1]
    if fc1.out_features != fc2.in_features:
        return None
    was_training = expert.training
    try:
        expert.eval()
        with torch.no_grad():
            x = torch.randn(4, fc1.in_features, dtype=fc1.weight.dtype, device=fc1.weight.device)
            y = expert(x)
            y = y[0] if isinstance(y, tuple) else y
            tol = 1e-5 if y.dtype == torch.float32 else 2e-2
            for name, act in _ACTIVATIONS:
                if y.shape == (4, fc2.out_features) and torch.allclose(fc2(act(fc1(x))).float(), y.float(), atol=tol, rtol=tol):
                    return name, act
    except Exception:
        return None
    finally:
        expert.train(was_training)
    return None


class Experts(torch.nn.Module):

    def __init__(self, expert, num_local_experts=1, expert_group_name=None):
        super().__init__()
        self.deepspeed_experts = torch.nn.ModuleList([copy.deepcopy(expert) for _ in range(num_local_experts)])
        self.num_local_experts = num_local_experts
        # expert parameters are NOT all-reduced over the dense data-parallel group: they belong to the expert-dp group
        for expert in self.deepspeed_experts:
            for _, param in expert.named_parameters():
                param.allreduce = False
                param.group_name = expert_group_name
        self._grouped_act = None
        if os.environ.get("DSB200_GROUPED_EXPERTS", "1") != "0":
            probe = _probe_two_layer_mlp(self.deepspeed_experts[0])
            if probe is not None:
                self._grouped_act = probe  # (name, fn)

    def _layers(self):
        out = []
        for e in self.deepspeed_experts:
            linears = [m for m in e.modules() if isinstance(m, torch.nn.Linear)]
            fc1, fc2 = linears if linears[0].out_features == linears[1].in_features else linears[::-1]
            out.append((fc1, fc2))
        return out

    def _grouped_ready(self, layers):
        """Weights must be whole (ZeRO-3 shards them between forwards and gathers through module hooks the grouped path skips)."""
        return all(l.weight.numel() == l.out_features * l.in_features and not l.weight.is_meta for pair in layers for l in pair)

    def _forward_grouped(self, inputs, layers):
        ep, E, C, M = inputs.shape
        x3 = inputs.permute(1, 0, 2, 3).reshape(E, ep * C, M)
        stack = lambda ts: torch.stack(list(ts)) if ts[0] is not None else None
        w1, b1 = stack([f.weight for f, _ in layers]), stack([f.bias for f, _ in layers])
        w2, b2 = stack([g.weight for _, g in layers]), stack([g.bias for _, g in layers])
        h = self._grouped_act[1](_GroupedLinearFn.apply(x3, w1, b1))
        y = _GroupedLinearFn.apply(h, w2, b2)
        return y.view(E, ep, C, -1).permute(1, 0, 2, 3)

    def forward(self, inputs):
        """inputs: [ep, local_E, C, M] -> same shape; expert e consumes chunk e along dim 1."""
        if self._grouped_act is not None and inputs.dim() == 4:
            layers = self._layers()
            if self._grouped_ready(layers):
                return self._forward_grouped(inputs, layers)
        chunks = inputs.chunk(self.num_local_experts, dim=1)
        expert_outputs = []
        for chunk, expert in zip(chunks, self.deepspeed_experts):
            out = expert(chunk)
            if isinstance(out, tuple):
                out = out[0]  # ignore bias terms of Megatron-style experts
            expert_outputs.append(out)
        return torch.cat(expert_outputs, dim=1)
