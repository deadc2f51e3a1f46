"""Token drop/gather across the tensor-parallel group for MoE with non-TP experts (parity: deepspeed/moe/mappings.py)."""
import torch

from .. import comm as dist
from ..utils import groups


def _gather_tokens(input_, dim=0):
    mpg = groups._get_model_parallel_group()
    world = dist.get_world_size(mpg)
    if world == 1:
        return input_
    input_ = input_.contiguous()
    tensor_list = [torch.empty_like(input_) for _ in range(world)]
    dist.all_gather(tensor_list, input_, group=mpg)
    return torch.cat(tensor_list, dim=dim).contiguous()


def _drop_tokens(input_, dim=0):
    mpg = groups._get_model_parallel_group()
    world = dist.get_world_size(mpg)
    if world == 1:
        return input_
    rank = dist.get_rank(mpg)
    assert input_.shape[dim] % world == 0, f"input dimension {dim} ({input_.shape[dim]}) is not divisible by tensor parallel world size ({world})"
    chunk = input_.shape[dim] // world
    return torch.narrow(input_, dim, rank * chunk, chunk)


class _GatherTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, dim):
        ctx.dim = dim
        return _gather_tokens(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _drop_tokens(grad_output, ctx.dim), None


class _DropTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, dim):
        ctx.dim = dim
        return _drop_tokens(input_, dim)

    @staticmethod
    def backward(ctx, input_):
        return _gather_tokens(input_, ctx.dim), None


def gather_tokens(input_, dim=0):
    if groups._get_model_parallel_world_size() == 1:
        return input_
    return _GatherTokens.apply(input_, dim)


def drop_tokens(input_, dim=0):
    if groups._get_model_parallel_world_size() == 1:
        return input_
    return _DropTokens.apply(input_, dim)
