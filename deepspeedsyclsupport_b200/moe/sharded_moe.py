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


def _sampler(cache, device, make):
    """Per-device random sampler, built once (the three module-level maps keep the reference's names)."""
    fn = cache.get(device)
    if fn is None:
        fn = cache[device] = make()
    return fn


def _uniform(lo, hi, device):
    return torch.distributions.uniform.Uniform(low=torch.tensor(lo, device=device), high=torch.tensor(hi, device=device)).rsample


def multiplicative_jitter(x, device, epsilon=1e-2):
    """x * U(1 - eps, 1 + eps), element-wise (the 'Jitter' noisy-gate policy)."""
    if epsilon == 0:
        return x
    return x * _sampler(uniform_map, device, lambda: _uniform(1.0 - epsilon, 1.0 + epsilon, device))(x.shape)


def gumbel_rsample(shape, device):
    make = lambda: torch.distributions.gumbel.Gumbel(torch.tensor(0.0, device=device), torch.tensor(1.0, device=device)).rsample
    return _sampler(gumbel_map, device, make)(shape)


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
    tokens, experts = gates.shape
    return max(math.ceil(tokens / experts * capacity_factor), min_capacity)


def _slot_positions(mask):
    """mask: [S, E] 0/1 -> position of every token inside its expert queue (exclusive cumsum along tokens)."""
    return torch.cumsum(mask, dim=0) - 1


def _load_balance(gates, first_choice_mask):
    """sum_e mean_s(gate[s, e]) * mean_s(first choice is e): the GShard auxiliary term before its normalisation constant."""
    return gates.mean(dim=0) * first_choice_mask.float().mean(dim=0)


def _no_drop_capacity(per_expert_load, device):
    """Largest queue over all experts and all ranks (the only host synchronisation of the gate, no-drop mode only)."""
    cap = per_expert_load.max().to(device)
    if dist.is_initialized():
        dist.all_reduce(cap, op=dist.ReduceOp.MAX, group=dist.get_world_group())
    return int(cap.item())


def _routes(mask, choice, slots):
    """(expert or -1 when dropped, slot) per token from a 0/1 [S, E] mask of KEPT assignments."""
    kept = mask.sum(dim=1) > 0
    return torch.where(kept, choice, torch.full_like(choice, -1)), (slots * mask).sum(dim=1)


def top1gating(logits, capacity_factor, min_capacity, used_token=None, noisy_gate_policy=None, drop_tokens=True,
               use_rts=True, use_tutel=False):
    """Returns (l_aux, token_expert [S,1], token_slot [S,1], token_weight [S,1], capacity, exp_counts)."""
    gates = F.softmax(logits, dim=1)
    experts = gates.shape[1]
    capacity = _capacity(gates, capacity_factor, min_capacity)
    ranked = logits + gumbel_rsample(logits.shape, device=logits.device) if noisy_gate_policy == "RSample" else gates
    choice = ranked.argmax(dim=1)
    mask = F.one_hot(choice, num_classes=experts)
    if used_token is not None:
        mask = used_token[:, None] * mask
    exp_counts = mask.sum(dim=0).detach()  # stays on the device: no host synchronisation in the gate (call .cpu() to log it)
    if not drop_tokens:
        capacity = _no_drop_capacity(exp_counts, logits.device)
    l_aux = _load_balance(gates, mask).sum() * experts
    assert logits.shape[0] >= min_capacity, \
        "No. of tokens (batch-size) should be greater than min_capacity. Either set min_capacity to 0 or increase your batch size."
    # every expert keeps `capacity` of its tokens: the ones with the highest priority -- a uniform draw per assignment under
    # random token selection (RTS), otherwise ties of 1s, i.e. whatever topk picks first
    priority = mask
    if use_rts:
        priority = mask * _sampler(exp_selection_uniform_map, logits.device, lambda: _uniform(0.0, 1.0, logits.device))(mask.shape)
    winners = torch.topk(priority, k=min(capacity, mask.shape[0]), dim=0).indices
    mask = mask * torch.zeros_like(mask).scatter_(0, winners, 1)
    token_expert, token_slot = _routes(mask, choice, _slot_positions(mask))
    weight = (gates * mask.float()).sum(dim=1)
    return l_aux, token_expert[:, None], token_slot[:, None], weight[:, None], capacity, exp_counts


def top2gating(logits, capacity_factor, min_capacity, drop_tokens=True):
    """Second expert = argmax of the gumbel-perturbed logits with the first one excluded; the two gate values are renormalised
    to sum to 1; the second choices queue behind ALL first choices of their expert."""
    gates = F.softmax(logits, dim=1)
    experts = gates.shape[1]
    capacity = _capacity(gates, capacity_factor * 2, min_capacity)
    first = gates.argmax(dim=1)
    m1 = F.one_hot(first, num_classes=experts)
    noisy = logits + gumbel_rsample(logits.shape, device=logits.device)
    second = noisy.masked_fill(m1.bool(), float("-inf")).argmax(dim=1)
    m2 = F.one_hot(second, num_classes=experts)
    pos1 = _slot_positions(m1)
    pos2 = _slot_positions(m2) + m1.sum(dim=0, keepdim=True)
    exp_counts = m1.sum(dim=0).detach()
    l_aux = _load_balance(gates, m1).mean() * experts * experts
    if drop_tokens:
        m1, m2 = m1 * (pos1 < capacity), m2 * (pos2 < capacity)
    else:
        capacity = _no_drop_capacity(exp_counts + m2.sum(dim=0).detach(), logits.device)
    e1, slot1 = _routes(m1, first, pos1)
    e2, slot2 = _routes(m2, second, pos2)
    g1, g2 = (gates * m1.float()).sum(dim=1), (gates * m2.float()).sum(dim=1)
    norm = (g1 + g2).clamp(min=torch.finfo(g1.dtype).eps)
    return (l_aux, torch.stack([e1, e2], dim=1), torch.stack([slot1, slot2], dim=1), torch.stack([g1 / norm, g2 / norm], dim=1), capacity,
            exp_counts)


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
