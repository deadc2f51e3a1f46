"""Random-LTD token sampling / gather / scatter (parity: deepspeed/ops/random_ltd/dropping_utils.py:15-133) on the
csrc/random_ltd kernels; CPU tensors take the plain torch path."""
import torch

from .. import native

native.register_sigs({
    "dsb200_token_sort": "p iii p",
    "dsb200_token_gather": "ppp iii l i p",
    "dsb200_token_scatter": "ppp iii l i p",
})


def token_sort_(sampled_indices, seq_length):
    """Sort every [..., reserved] row of an int32 index tensor ascending, in place."""
    assert sampled_indices.dtype == torch.int32 and sampled_indices.is_contiguous()
    if not sampled_indices.is_cuda:
        sampled_indices.copy_(sampled_indices.sort(dim=-1)[0])
        return sampled_indices
    reserved = sampled_indices.shape[-1]
    rows = sampled_indices.numel() // reserved
    native.check(native.lib().dsb200_token_sort(native.ptr(sampled_indices), rows, reserved, int(seq_length), native.stream_ptr()))
    return sampled_indices


def _dims(activations, sorted_indices, batch_first):
    if batch_first:
        B, S = activations.shape[0], activations.shape[1]
    else:
        S, B = activations.shape[0], activations.shape[1]
    R = sorted_indices.shape[-1]
    row_bytes = activations[0, 0].numel() * activations.element_size()
    return B, S, R, row_bytes


def token_gather(activations, sorted_indices, batch_first):
    B, S, R, row_bytes = _dims(activations, sorted_indices, batch_first)
    if not activations.is_cuda:
        ix = sorted_indices.long()
        if batch_first:
            return torch.gather(activations, 1, ix.view(B, R, *([1] * (activations.dim() - 2))).expand(B, R, *activations.shape[2:]))
        return torch.gather(activations, 0, ix.t().reshape(R, B, *([1] * (activations.dim() - 2))).expand(R, B, *activations.shape[2:]))
    activations = activations.contiguous()
    sorted_indices = sorted_indices.contiguous()
    shape = (B, R, *activations.shape[2:]) if batch_first else (R, B, *activations.shape[2:])
    out = torch.empty(shape, dtype=activations.dtype, device=activations.device)
    native.check(native.lib().dsb200_token_gather(native.ptr(activations), native.ptr(out), native.ptr(sorted_indices), B, S, R, row_bytes,
                                                  int(batch_first), native.stream_ptr()))
    return out


def token_scatter_(all_activations, layer_activations, sorted_indices, batch_first):
    B, S, R, row_bytes = _dims(all_activations, sorted_indices, batch_first)
    if not all_activations.is_cuda:
        ix = sorted_indices.long()
        if batch_first:
            all_activations.scatter_(1, ix.view(B, R, *([1] * (all_activations.dim() - 2))).expand_as(layer_activations), layer_activations)
        else:
            all_activations.scatter_(0, ix.t().reshape(R, B, *([1] * (all_activations.dim() - 2))).expand_as(layer_activations), layer_activations)
        return all_activations
    assert all_activations.is_contiguous()
    layer_activations = layer_activations.contiguous()
    native.check(native.lib().dsb200_token_scatter(native.ptr(all_activations), native.ptr(layer_activations), native.ptr(sorted_indices.contiguous()),
                                                   B, S, R, row_bytes, int(batch_first), native.stream_ptr()))
    return all_activations


def _sample(reserved_length, seq_length, batch_size, layers, device):
    prob_dist = torch.ones((layers * batch_size, seq_length), device=device)
    sampled_indices = torch.multinomial(prob_dist, reserved_length)
    sampled_indices = sampled_indices.reshape(layers, batch_size, reserved_length).to(torch.int32).contiguous()
    return token_sort_(sampled_indices, seq_length)


def gpt_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1, device: str = "cpu", attn_mask: torch.Tensor = None):
    """Returns sampled_indices [layers, batch, reserved] and the causal mask cropped to [batch, 1, reserved, reserved]."""
    sampled_indices = _sample(reserved_length, seq_length, batch_size, layers, device)
    new_mask = attn_mask[:, :, :reserved_length, :reserved_length] if attn_mask is not None else None
    return sampled_indices, new_mask


def bert_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1, device: str = "cpu", attn_mask: torch.Tensor = None):
    """Returns sampled_indices [layers, batch, reserved] and per-layer masks [batch, 1, reserved, reserved] gathered on both axes."""
    assert attn_mask is not None
    sampled_indices = _sample(reserved_length, seq_length, batch_size, layers, device)
    ix = sampled_indices.long()
    new_mask = []
    for l in range(layers):
        rows = torch.gather(attn_mask, 2, ix[l].view(batch_size, 1, reserved_length, 1).expand(batch_size, attn_mask.shape[1], reserved_length, attn_mask.shape[3]))
        new_mask.append(torch.gather(rows, 3, ix[l].view(batch_size, 1, 1, reserved_length).expand(batch_size, attn_mask.shape[1], reserved_length, reserved_length)))
    return sampled_indices, new_mask


class GatherTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, activations: torch.Tensor, sorted_indices: torch.Tensor, batch_first: bool):
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return activations, token_gather(activations, sorted_indices, batch_first)

    @staticmethod
    def backward(ctx, a_gradients: torch.Tensor, g_gradients: torch.Tensor):
        sorted_indices, = ctx.saved_tensors
        # the kept tokens' gradient through the gathered copy overwrites (not adds to) the pass-through gradient at
        # those positions -- this mirrors upstream, where the pass-through branch is replaced by ScatterTokens later
        return token_scatter_(a_gradients.contiguous().clone(), g_gradients.contiguous(), sorted_indices, ctx.batch_first), None, None


class ScatterTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, all_activations: torch.Tensor, layer_activations: torch.Tensor, sorted_indices: torch.Tensor, batch_first: bool):
        scatter_results = token_scatter_(all_activations.contiguous().clone(), layer_activations, sorted_indices, batch_first)
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return scatter_results

    @staticmethod
    def backward(ctx, out_gradients: torch.Tensor):
        out_gradients = out_gradients.contiguous()
        sorted_indices, = ctx.saved_tensors
        ret_val = token_gather(out_gradients, sorted_indices, ctx.batch_first)
        return out_gradients, ret_val, None, None
