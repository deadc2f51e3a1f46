"""GShard-style top-1 / top-2 gated mixture of experts with expert parallelism.

Semantics follow deepspeed/moe/sharded_moe.py: capacity = ceil(S/E * cf) (:162-171), top-1 with optional random
token selection (RTS) and noisy gating (:184-279), top-2 with gumbel second choice and weight normalisation
(:282-345), load-balancing aux loss, drop tokens over capacity or no-drop with all-reduced max capacity (:213-217).

Mechanics are index-based instead of the reference's dense one-hot einsums ("sec,sm->ecm" :500, "sec,ecm->sm" :546):
each token carries (expert, slot, weight); dispatch is a row scatter into the [E, C, M] buffer, combine is a weighted
row gather -- O(S*k*M) instead of O(S*E*C*M).  On CUDA the gate / scatter / gather run as hand-written kernels
(csrc/moe), and the expert-parallel exchange is `all_to_all_single` (NCCL baseline) or the fused peer-memory
dispatch in parallel/fused.
"""
import math
from typing import Callable, Dict, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor
from torch.nn import Module

from .. import comm as dist
from ..utils import groups
from ..utils.timer import SynchronizedWallClockTimer

uniform_map: Dict[torch.device, Callable] = {}
gumbel_map: Dict[torch.device, Callable] = {}
exp_selection_uniform_map: Dict[torch.device, Callable] = {}

TOPK_GATE_TIMER = "topk_gate"
MOE_TIMER = "moe"
FIRST_ALLTOALL_TIMER = "1st_a2a"
SECOND_ALLTOALL_TIMER = "2nd_a2a"


def multiplicative_jitter(x, device, epsilon=1e-2):
    if epsilon == 0:
        return x
    uniform = uniform_map.get(device)
    if uniform is None:
        uniform = torch.distributions.uniform.Uniform(low=torch.tensor(1.0 - epsilon, device=device),
                                                      high=torch.tensor(1.0 + epsilon, device=device)).rsample
        uniform_map[device] = uniform
    return x * uniform(x.shape)


def gumbel_rsample(shape, device):
    gumbel = gumbel_map.get(device)
    if gumbel is None:
        one = torch.tensor(1.0, device=device)
        zero = torch.tensor(0.0, device=device)
        gumbel = torch.distributions.gumbel.Gumbel(zero, one).rsample
        gumbel_map[device] = gumbel
    return gumbel(shape)


class _AllToAll(torch.autograd.Function):

    @staticmethod
    def forward(ctx, group, input):
        ctx.group = group
        input = input.contiguous()
        output = torch.empty_like(input)
        dist.all_to_all_single(output, input, group=group)
        return output

    @staticmethod
    def backward(ctx, *grad_output):
        return (None, _AllToAll.apply(ctx.group, *grad_output))


def _capacity(gates, capacity_factor, min_capacity):
    num_tokens, num_experts = gates.shape[0], gates.shape[1]
    capacity = math.ceil((num_tokens / num_experts) * capacity_factor)
    return max(capacity, min_capacity)


def _slot_positions(mask):
    """mask: [S, E] 0/1 -> position of every token inside its expert queue (exclusive cumsum along tokens)."""
    return torch.cumsum(mask, dim=0) - 1


def top1gating(logits, capacity_factor, min_capacity, used_token=None, noisy_gate_policy=None, drop_tokens=True,
               use_rts=True, use_tutel=False):
    """Returns (l_aux, token_expert [S,1], token_slot [S,1], token_weight [S,1], capacity, exp_counts)."""
    if noisy_gate_policy == "RSample":
        logits_w_noise = logits + gumbel_rsample(logits.shape, device=logits.device)
    gates = F.softmax(logits, dim=1)
    capacity = _capacity(gates, capacity_factor, min_capacity)
    indices1_s = torch.argmax(logits_w_noise if noisy_gate_policy == "RSample" else gates, dim=1)
    num_experts = int(gates.shape[1])
    mask1 = F.one_hot(indices1_s, num_classes=num_experts)
    if used_token is not None:
        mask1 = torch.einsum("s,se->se", used_token, mask1)
    exp_counts = torch.sum(mask1, dim=0).detach()  # stays on the device: no host synchronisation in the gate (call .cpu() to log it)
    if not drop_tokens:
        new_capacity = torch.max(exp_counts).to(logits.device)
        if dist.is_initialized():
            dist.all_reduce(new_capacity, op=dist.ReduceOp.MAX, group=dist.get_world_group())
        capacity = int(new_capacity.item())
    me = torch.mean(gates, dim=0)
    ce = torch.mean(mask1.float(), dim=0)
    l_aux = torch.sum(me * ce) * num_experts
    if use_rts:
        uniform = exp_selection_uniform_map.get(logits.device)
        if uniform is None:
            uniform = torch.distributions.uniform.Uniform(low=torch.tensor(0.0, device=logits.device),
                                                          high=torch.tensor(1.0, device=logits.device)).rsample
            exp_selection_uniform_map[logits.device] = uniform
        mask1_rand = mask1 * uniform(mask1.shape)
    else:
        mask1_rand = mask1
    assert logits.shape[0] >= min_capacity, \
        "No. of tokens (batch-size) should be greater than min_capacity. Either set min_capacity to 0 or increase your batch size."
    # keep the top-`capacity` tokens of every expert (by random priority under RTS, by arrival order otherwise)
    k = min(capacity, mask1_rand.shape[0])
    top_idx = torch.topk(mask1_rand, k=k, dim=0)[1]
    new_mask1 = mask1 * torch.zeros_like(mask1).scatter_(0, top_idx, 1)
    mask1 = new_mask1
    locations1 = _slot_positions(mask1)
    kept = mask1.sum(dim=1) > 0
    token_slot = torch.sum(locations1 * mask1, dim=1)
    gate_w = (gates * mask1.float()).sum(dim=1)
    token_expert = torch.where(kept, indices1_s, torch.full_like(indices1_s, -1))
    return l_aux, token_expert[:, None], token_slot[:, None], gate_w[:, None], capacity, exp_counts


def top2gating(logits, capacity_factor, min_capacity, drop_tokens=True):
    gates = F.softmax(logits, dim=1)
    capacity = _capacity(gates, capacity_factor * 2, min_capacity)
    indices1_s = torch.argmax(gates, dim=1)
    num_experts = int(gates.shape[1])
    mask1 = F.one_hot(indices1_s, num_classes=num_experts)
    logits_w_noise = logits + gumbel_rsample(logits.shape, device=logits.device)
    logits_except1 = logits_w_noise.masked_fill(mask1.bool(), float("-inf"))
    indices2_s = torch.argmax(logits_except1, dim=1)
    mask2 = F.one_hot(indices2_s, num_classes=num_experts)
    locations1 = _slot_positions(mask1)
    locations2 = _slot_positions(mask2) + torch.sum(mask1, dim=0, keepdim=True)
    exp_counts = torch.sum(mask1, dim=0).detach()
    me = torch.mean(gates, dim=0)
    ce = torch.mean(mask1.float(), dim=0)
    l_aux = torch.mean(me * ce) * num_experts * num_experts
    if drop_tokens:
        mask1 = mask1 * torch.lt(locations1, capacity)
        mask2 = mask2 * torch.lt(locations2, capacity)
    else:
        new_capacity = torch.max(exp_counts + torch.sum(mask2, dim=0).detach()).to(logits.device)
        if dist.is_initialized():
            dist.all_reduce(new_capacity, op=dist.ReduceOp.MAX, group=dist.get_world_group())
        capacity = int(new_capacity.item())
    slot1 = torch.sum(locations1 * mask1, dim=1)
    slot2 = torch.sum(locations2 * mask2, dim=1)
    mask1_float, mask2_float = mask1.float(), mask2.float()
    gates1_s = torch.einsum("se,se->s", gates, mask1_float)
    gates2_s = torch.einsum("se,se->s", gates, mask2_float)
    denom_s = torch.clamp(gates1_s + gates2_s, min=torch.finfo(gates1_s.dtype).eps)
    gates1_s = gates1_s / denom_s
    gates2_s = gates2_s / denom_s
    kept1, kept2 = mask1.sum(1) > 0, mask2.sum(1) > 0
    e1 = torch.where(kept1, indices1_s, torch.full_like(indices1_s, -1))
    e2 = torch.where(kept2, indices2_s, torch.full_like(indices2_s, -1))
    token_expert = torch.stack([e1, e2], dim=1)
    token_slot = torch.stack([slot1, slot2], dim=1)
    token_weight = torch.stack([gates1_s, gates2_s], dim=1)
    return l_aux, token_expert, token_slot, token_weight, capacity, exp_counts


def dense_dispatch_masks(token_expert, token_slot, token_weight, num_experts, capacity):
    """(combine_weights [S,E,C], dispatch_mask [S,E,C]) in the reference's dense form -- used by tests / oracles."""
    S, K = token_expert.shape
    cw = torch.zeros(S, num_experts, capacity, dtype=token_weight.dtype, device=token_weight.device)
    for k in range(K):
        valid = token_expert[:, k] >= 0
        idx = torch.nonzero(valid).squeeze(1)
        cw[idx, token_expert[idx, k], token_slot[idx, k]] += token_weight[idx, k]
    return cw, cw.bool()


class TopKGate(Module):
    """Gate module: fp32 linear router + top-k assignment (parity: sharded_moe.py:348-422)."""

    def __init__(self, model_dim, num_experts, k=1, capacity_factor=1.0, eval_capacity_factor=1.0, min_capacity=8,
                 noisy_gate_policy=None, drop_tokens=True, use_rts=True):
        super().__init__()
        if k != 1 and k != 2:
            raise ValueError("Only top-1 and top-2 gatings are supported.")
        self.wg = torch.nn.Linear(model_dim, num_experts, bias=False).float()
        self.k = k
        self.capacity_factor = capacity_factor
        self.eval_capacity_factor = eval_capacity_factor
        self.min_capacity = min_capacity
        self.noisy_gate_policy = noisy_gate_policy
        self.timers = SynchronizedWallClockTimer()
        self.wall_clock_breakdown = False
        self.gate_time = 0.0
        self.drop_tokens = drop_tokens
        self.use_rts = use_rts

    def forward(self, input, used_token=None, use_tutel=False):
        if self.wall_clock_breakdown:
            self.timers(TOPK_GATE_TIMER).start()
        if self.wg.weight.dtype != torch.float32:
            self.wg = self.wg.float()
        input_fp32 = input.float()
        if self.noisy_gate_policy == "Jitter" and self.training:
            input_fp32 = multiplicative_jitter(input_fp32, device=input.device)
        logits = F.linear(input_fp32, self.wg.weight.float())
        if self.k == 1:
            out = top1gating(logits, self.capacity_factor if self.training else self.eval_capacity_factor, self.min_capacity,
                             used_token, self.noisy_gate_policy if self.training else None, self.drop_tokens, self.use_rts)
        else:
            out = top2gating(logits, self.capacity_factor if self.training else self.eval_capacity_factor, self.min_capacity,
                             self.drop_tokens)
        if self.wall_clock_breakdown:
            self.timers(TOPK_GATE_TIMER).stop()
            self.gate_time = self.timers(TOPK_GATE_TIMER).elapsed(reset=False)
        return out


def dispatch_tokens(x, token_expert, token_slot, num_experts, capacity):
    """x [S, M] -> [E, C, M]; differentiable.  Dropped tokens (expert < 0) go to a dummy row that is cut off again, so no
    data-dependent shape (and no host synchronisation) is involved."""
    S, M = x.shape
    K = token_expert.shape[1]
    rows = num_experts * capacity
    dst = torch.where(token_expert >= 0, token_expert * capacity + token_slot, torch.full_like(token_expert, rows))
    src = x if K == 1 else x.repeat_interleave(K, dim=0)
    out = x.new_zeros(rows + 1, M).index_add(0, dst.reshape(-1), src)
    return out[:rows].view(num_experts, capacity, M)


def combine_tokens(expert_out, token_expert, token_slot, token_weight, num_tokens):
    """expert_out [E, C, M] -> [S, M] = sum_k w_k * expert_out[e_k, slot_k] (dropped tokens read a zero row)."""
    E, C, M = expert_out.shape
    K = token_expert.shape[1]
    flat = torch.cat([expert_out.reshape(E * C, M), expert_out.new_zeros(1, M)])
    valid = token_expert >= 0
    src = torch.where(valid, token_expert * C + token_slot, torch.full_like(token_expert, E * C))
    rows = flat.index_select(0, src.reshape(-1)).view(num_tokens, K, M)
    w = (token_weight * valid.to(token_weight.dtype)).to(flat.dtype)
    return (rows * w[:, :, None]).sum(dim=1)


class MOELayer(Module):
    """Gate -> dispatch -> all-to-all -> local experts -> all-to-all -> combine (parity: sharded_moe.py:425-554)."""

    def __init__(self, gate, experts, ep_group_name, ep_size, num_local_experts, use_tutel=False):
        super().__init__()
        self.gate = gate
        self.experts = experts
        self.ep_group = None
        self.ep_size = ep_size
        self.ep_group_name = ep_group_name
        self.num_local_experts = num_local_experts
        self.time_falltoall = 0.0
        self.time_salltoall = 0.0
        self.time_moe = 0.0
        self.timers = SynchronizedWallClockTimer()
        self.wall_clock_breakdown = False
        self.use_tutel = False

    def _set_ep_group(self, ep_group):
        self.ep_group = ep_group

    def _forward_peer(self, original, tokens, token_expert, token_slot, token_weight, capacity):
        """Expert parallelism inside one NVSwitch domain: dispatch = one push kernel into the owners' receive buffers,
        combine = one weighted pull kernel (moe/peer_dispatch.py, csrc/comm/symm.cu); no all-to-all, no scatter tensor."""
        from .peer_dispatch import PeerCombine, PeerDispatch, PeerMoEBuffers
        d_model = tokens.shape[-1]
        bufs = PeerMoEBuffers.get(self.ep_group, self.num_local_experts, int(capacity), d_model)
        received = PeerDispatch.apply(tokens, token_expert, token_slot, bufs)             # [ep, n_local, C_buf, M]
        expert_output = self.experts(received)
        out = PeerCombine.apply(expert_output.contiguous(), token_expert, token_slot, token_weight, tokens.shape[0], bufs)
        if self.wall_clock_breakdown:
            self.timers(MOE_TIMER).stop()
            self.time_moe = self.timers(MOE_TIMER).elapsed(reset=False)
        return out.reshape(original.shape)

    def forward(self, *input, **kwargs):
        if self.wall_clock_breakdown:
            self.timers(MOE_TIMER).start()
        d_model = input[0].shape[-1]
        reshaped_input = input[0].reshape(-1, d_model)
        self.l_aux, token_expert, token_slot, token_weight, capacity, self.exp_counts = self.gate(
            reshaped_input, input[1] if len(input) > 1 else None)
        E = self.ep_size * self.num_local_experts
        from . import peer_dispatch
        if peer_dispatch.usable(reshaped_input, self.ep_size, self.ep_group) and not (
                groups._get_expert_model_parallel_world_size() == 1 and groups._get_model_parallel_world_size() > 1):
            return self._forward_peer(input[0], reshaped_input, token_expert, token_slot, token_weight, capacity)
        dispatched_input = dispatch_tokens(reshaped_input, token_expert, token_slot, E, capacity)
        if self.wall_clock_breakdown:
            self.timers(FIRST_ALLTOALL_TIMER).start()
        if groups._get_expert_model_parallel_world_size() == 1 and groups._get_model_parallel_world_size() > 1:
            from .mappings import drop_tokens
            dispatched_input = drop_tokens(dispatched_input, dim=1)
        if self.ep_size > 1:
            dispatched_input = _AllToAll.apply(self.ep_group, dispatched_input)
        if self.wall_clock_breakdown:
            self.timers(FIRST_ALLTOALL_TIMER).stop()
            self.time_falltoall = self.timers(FIRST_ALLTOALL_TIMER).elapsed(reset=False)
        dispatched_input = dispatched_input.reshape(self.ep_size, self.num_local_experts, -1, d_model)
        expert_output = self.experts(dispatched_input)
        if self.wall_clock_breakdown:
            self.timers(SECOND_ALLTOALL_TIMER).start()
        if self.ep_size > 1:
            expert_output = _AllToAll.apply(self.ep_group, expert_output.contiguous())
        if self.wall_clock_breakdown:
            self.timers(SECOND_ALLTOALL_TIMER).stop()
            self.time_salltoall = self.timers(SECOND_ALLTOALL_TIMER).elapsed(reset=False)
        expert_output = expert_output.reshape(E, -1, d_model)
        if groups._get_expert_model_parallel_world_size() == 1 and groups._get_model_parallel_world_size() > 1:
            from .mappings import gather_tokens
            expert_output = gather_tokens(expert_output, dim=1)
        combined_output = combine_tokens(expert_output, token_expert, token_slot, token_weight, reshaped_input.shape[0])
        a = combined_output.reshape(input[0].shape)
        if self.wall_clock_breakdown:
            self.timers(MOE_TIMER).stop()
            self.time_moe = self.timers(MOE_TIMER).elapsed(reset=False)
        return a
