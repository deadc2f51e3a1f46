"""Token drop / gather across the tensor-parallel group for MoE with non-TP experts (parity: deepspeed/moe/mappings.py).

With tensor-parallel attention but replicated (non-TP) experts every TP rank holds the same tokens; `drop_tokens` keeps this rank's
1/tp slice along `dim` before the MoE layer and `gather_tokens` re-assembles the full tensor after it.  The two are each other's
adjoint, so ONE autograd function parametrised by the direction serves both."""
import torch

from .. import comm as dist
from ..utils import groups


def _tp():
    group = groups._get_model_parallel_group()
    return group, dist.get_world_size(group)


def _gather_tokens(input_, dim=0):
    group, world = _tp()
    if world == 1:
        return input_
    src = input_.contiguous()
    pieces = [torch.empty_like(src) for _ in range(world)]
    dist.all_gather(pieces, src, group=group)
    return torch.cat(pieces, dim=dim).contiguous()


def _drop_tokens(input_, dim=0):
    group, world = _tp()
    if world == 1:
        return input_
    size = input_.shape[dim]
    assert size % world == 0, f"input dimension {dim} ({size}) is not divisible by tensor parallel world size ({world})"
    return input_.narrow(dim, dist.get_rank(group) * (size // world), size // world)


class _Resplit(torch.autograd.Function):
    """forward = gather (or drop), backward = the other one"""

    @staticmethod
    def forward(ctx, input_, dim, gather):
        ctx.dim, ctx.gather = dim, gather
        return (_gather_tokens if gather else _drop_tokens)(input_, dim)

    @staticmethod
    def backward(ctx, grad):
        return (_drop_tokens if ctx.gather else _gather_tokens)(grad, ctx.dim), None, None


# reference names of the two directions
class _GatherTokens(_Resplit):

    @staticmethod
    def forward(ctx, input_, dim):
        return _Resplit.forward(ctx, input_, dim, True)

    @staticmethod
    def backward(ctx, grad):
        return _Resplit.backward(ctx, grad)[:2]


class _DropTokens(_Resplit):

    @staticmethod
    def forward(ctx, input_, dim):
        return _Resplit.forward(ctx, input_, dim, False)

    @staticmethod
    def backward(ctx, grad):
        return _Resplit.backward(ctx, grad)[:2]


def gather_tokens(input_, dim=0):
    return input_ if groups._get_model_parallel_world_size() == 1 else _GatherTokens.apply(input_, dim)


def drop_tokens(input_, dim=0):
    return input_ if groups._get_model_parallel_world_size() == 1 else _DropTokens.apply(input_, dim)
