"""Pure-torch random token drop for [batch, seq, hidden] decoder activations, batched (no per-sample Python loop):
a random permutation per row comes from argsort of uniform noise, the first `reserved_length` entries are sorted to keep
causal order.  Same call signatures as the reference helpers (data_routing/utils.py)."""
import torch


def bsh_decoder_gather(reserved_length, hidden_states, mask):
    B, S, H = hidden_states.shape
    noise = torch.rand(B, S, device=hidden_states.device)
    keep = noise.argsort(dim=1)[:, :reserved_length].sort(dim=1).values  # [B, R], ascending positions
    part_hidden_states = torch.gather(hidden_states, 1, keep.unsqueeze(-1).expand(B, reserved_length, H))
    part_mask = mask[:, :, :reserved_length, :reserved_length]
    return part_hidden_states, list(keep.unbind(0)), part_mask


def bsh_decoder_scatter(hidden_states, part_hidden_states, rand_list):
    keep = torch.stack(list(rand_list), 0)
    hidden_states.scatter_(1, keep.unsqueeze(-1).expand_as(part_hidden_states), part_hidden_states)
    return hidden_states
