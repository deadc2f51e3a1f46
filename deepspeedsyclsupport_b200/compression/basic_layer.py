"""Compression-aware layers.

One mixin (`_CompressState`) owns everything the techniques share -- pruning masks (l1: fixed buffer computed at
enable time; topk: learnable scores binarised with a straight-through top-k), the weight / activation quantizer choice
and the "fix" operations that bake a technique into the weights (optionally shrinking the layer and returning the mask so
the FOLLOWING layer can drop the matching input columns).  Linear / Conv2d / Embedding / BatchNorm subclasses only
describe how a mask maps onto their weight layout.  The attribute names toggled by the scheduler
(`weight_quantization_enabled`, `sparse_pruning_enabled`, ...) and the enable_* / fix_* method names follow the
reference (deepspeed/compression/basic_layer.py) since user code and the scheduler address them by name."""
import math

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn import init

from .. import comm as dist
from ..utils.logging import logger
from .utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer, TopKBinarizer

g_mpu = None


def _pick_weight_quantizer(target_bits, quantization_type):
    if target_bits >= 3:
        return SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply
    assert quantization_type == "symmetric", "Only symmetric quantization is supported for ternary/binary weight quantization"
    return TernaryQuantizer.apply if target_bits == 2 else BinaryQuantizer.apply


class QuantAct(nn.Module):
    """Static-range activation quantiser: tracks a running min/max with momentum during training and quantises with it."""

    def __init__(self, act_range_momentum=0.95, quant_mode="symmetric"):
        super().__init__()
        self.act_range_momentum = act_range_momentum
        self.quant_mode = quant_mode
        self.act_function = SymQuantizer.apply if quant_mode == "symmetric" else AsymQuantizer.apply
        self.register_buffer("x_min_max", torch.zeros(2))

    def forward(self, x, num_bits, *args):
        if self.training:
            lo, hi = x.detach().min(), x.detach().max()
            if self.x_min_max[0] == self.x_min_max[1]:
                self.x_min_max[0], self.x_min_max[1] = lo, hi
            m = self.act_range_momentum
            self.x_min_max[0] = self.x_min_max[0] * m + lo * (1 - m)
            self.x_min_max[1] = self.x_min_max[1] * m + hi * (1 - m)
        return self.act_function(x, num_bits, self.x_min_max[0], self.x_min_max[1])


class _CompressState(object):
    """Shared technique state.  `_unit_shape(kind)` (per layer type) gives the shape of the score / mask tensor of a
    pruning kind; `_expand(kind, mask)` broadcasts a mask onto the weight."""

    def _init_compress_state(self):
        self.sparse_pruning_method = None
        self.row_pruning_method = None
        self.head_pruning_method = None
        self.channel_pruning_method = None
        self.activation_quantization_method = None
        self.weight.start_bits = None
        self.weight.target_bits = None
        self.weight.q_period = None
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False
        self.sparse_pruning_enabled = False
        self.row_pruning_enabled = False
        self.head_pruning_enabled = False
        self.channel_pruning_enabled = False
        self.activation_quantization_enabled = False

    # ---- helpers
    def _replace_weight(self, new_data):
        """Swap in a differently shaped weight but keep the MoQ annotations riding on the Parameter."""
        tags = (self.weight.start_bits, self.weight.target_bits, self.weight.q_period)
        self.weight = nn.Parameter(new_data)
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = tags

    def _enable_pruning(self, kind, ratio, method, norms_fn, score_shape, mask_view):
        setattr(self, f"{kind}_pruning_ratio", ratio)
        setattr(self, f"{kind}_pruning_method", method)
        scores_name = {"sparse": "sparse_mask_scores", "row": "row_mask_scores", "channel": "channel_mask_scores"}[kind]
        if method == "l1":
            mask = TopKBinarizer.apply(norms_fn(), ratio, False).view(*mask_view).to(self.weight.device)
        elif method == "topk":
            scores = nn.Parameter(torch.empty(*score_shape, device=self.weight.device))
            init.kaiming_uniform_(scores, a=math.sqrt(5))
            setattr(self, scores_name, scores)
            mask = None
        else:
            raise NotImplementedError
        self.register_buffer(f"{kind}_pruning_mask", mask)

    def get_mask(self, pruning_type="row"):
        if pruning_type == "head":
            if self.head_pruning_method != "topk":
                raise NotImplementedError
            return TopKBinarizer.apply(self.head_pruning_scores, self.head_pruning_ratio, False)
        if pruning_type not in ("sparse", "row", "channel"):
            raise NotImplementedError
        method = getattr(self, f"{pruning_type}_pruning_method")
        if method == "l1":
            return getattr(self, f"{pruning_type}_pruning_mask").to(self.weight.device)
        if method == "topk":
            scores = getattr(self, {"sparse": "sparse_mask_scores", "row": "row_mask_scores", "channel": "channel_mask_scores"}[pruning_type])
            return TopKBinarizer.apply(scores, getattr(self, f"{pruning_type}_pruning_ratio"), False)
        raise NotImplementedError

    def _drop_pruning_state(self, kind):
        scores_name = {"sparse": "sparse_mask_scores", "row": "row_mask_scores", "channel": "channel_mask_scores"}[kind]
        if hasattr(self, f"{kind}_pruning_mask"):
            delattr(self, f"{kind}_pruning_mask")
        if getattr(self, f"{kind}_pruning_method") == "topk" and hasattr(self, scores_name):
            delattr(self, scores_name)
        setattr(self, f"{kind}_pruning_method", None)
        setattr(self, f"{kind}_pruning_enabled", False)

    # ---- quantisation
    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits = start_bits
        self.weight.target_bits = target_bits
        self.weight.q_period = quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if weight_quantization_enabled_in_forward:
            logger.warning("************ A lot of MoQ features are not supported in quantize_weight_in_forward mode, please consider to use DS-FP16 optimizer************")
            self.weight_quantizer = _pick_weight_quantizer(target_bits, quantization_type)
            self.weight_quantize_num_groups = num_groups

    def fix_weight_quantization(self):
        self.weight.data = self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False
        return None

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        assert bits in [4, 8], "Only 4/8 bits activation quantization are supported for now"
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        if range_calibration == "static":
            self.activation_quantizer = QuantAct(quant_mode=quantization_type)
        else:
            self.activation_quantizer = SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply

    def _quantized_weight(self):
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            return self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups)
        return self.weight

    def _quantized_input(self, x):
        if not self.activation_quantization_enabled:
            return x
        groups = x.numel() // x.size(-1) if "dynamic" in self.activation_quantization_method else 1
        return self.activation_quantizer(x, self.activation_quantization_bits, None, None, groups)

    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.get_mask(pruning_type="sparse").view(self.weight.size())
        self._drop_pruning_state("sparse")
        return None


class LinearLayer_Compress(nn.Linear, _CompressState):
    """nn.Linear with weight/activation quantisation and sparse / row / head pruning."""

    def __init__(self, *kargs, bias=True):
        super().__init__(*kargs, bias=bias)
        self._init_compress_state()

    def extra_repr(self):
        return ("in_features={}, out_features={}, bias={}, sparse pruning={}, row pruning={}, head pruning={}, activation quantization={}, "
                "weight_quantization={}").format(self.in_features, self.out_features, self.bias is not None, self.sparse_pruning_method is not None,
                                                 self.row_pruning_method is not None, self.head_pruning_method is not None,
                                                 self.activation_quantization_method is not None, self.weight.target_bits)

    def enable_sparse_pruning(self, ratio, method):
        self._enable_pruning("sparse", ratio, method, lambda: self.weight.data.abs(), self.weight.size(), self.weight.size())

    def enable_row_pruning(self, ratio, method):
        self._enable_pruning("row", ratio, method, lambda: torch.linalg.norm(self.weight.data, ord=1, dim=1), (self.weight.size(0), 1), (-1, 1))

    def enable_head_pruning(self, ratio, method, num_heads):
        # applied to the attention OUTPUT projection: heads are groups of input columns
        if method not in ["topk"]:
            raise NotImplementedError
        self.num_heads = num_heads
        self.head_pruning_ratio = ratio
        self.head_pruning_method = method
        self.head_pruning_scores = nn.Parameter(torch.empty(1, num_heads, device=self.weight.device))
        init.kaiming_uniform_(self.head_pruning_scores, a=math.sqrt(5))

    def head_pruning_reshape(self, w, mask):
        out_f, in_f = w.shape
        return (w.t().reshape(self.num_heads, -1) * mask.view(-1, 1)).reshape(in_f, out_f).t()

    def fix_row_col_pruning_helper(self, mask=None, dim_reduction=False):
        """mask=None: bake MY row mask (zeroing rows, or removing them when dim_reduction) and return it;
        mask given: I am the layer that consumes a pruned layer's output -> drop the matching input columns."""
        if mask is not None:
            self._replace_weight(self.weight.data[:, mask.view(-1)])
            self.in_features = self.weight.size(1)
            self.row_pruning_enabled = False
            return None
        mask = self.get_mask(pruning_type="row").bool()
        keep = mask.view(-1)
        if dim_reduction:
            self._replace_weight(self.weight.data[keep, :])
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data[keep])
            self.out_features = self.weight.size(0)
        else:
            self.weight.data = self.weight.data * keep.view(-1, 1)
            if self.bias is not None:
                self.bias.data = self.bias.data * keep
        self._drop_pruning_state("row")
        return mask

    def fix_head_pruning_helper(self, mask=None, num_heads=None, dim_reduction=False):
        num_heads = num_heads if num_heads else self.num_heads
        if mask is not None:
            # a Q/K/V projection feeding the pruned output projection: heads are groups of output rows here
            in_f = self.weight.size(1)
            self._replace_weight(self.weight.data.reshape(num_heads, -1)[mask.view(-1), :].reshape(-1, in_f))
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data.reshape(num_heads, -1)[mask.view(-1), :].reshape(-1))
            self.head_pruning_enabled = False
            return mask
        if self.head_pruning_method != "topk":
            raise NotImplementedError
        mask = self.get_mask(pruning_type="head").bool()
        out_f = self.weight.size(0)
        if dim_reduction:
            self._replace_weight(self.weight.data.t().reshape(num_heads, -1)[mask.view(-1), :].reshape(-1, out_f).t())
        else:
            self.weight.data = self.head_pruning_reshape(self.weight.data, mask)
        del self.head_pruning_scores
        self.head_pruning_method = None
        self.head_pruning_enabled = False
        return mask

    def _effective(self):
        weight, bias = self._quantized_weight(), self.bias
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            weight = weight * self.get_mask(pruning_type="sparse").view(self.weight.size())
        if self.row_pruning_enabled and self.row_pruning_method:
            mask = self.get_mask(pruning_type="row")
            weight = weight * mask.view(-1, 1)
            if bias is not None:
                bias = bias * mask.view(-1)
        if self.head_pruning_enabled and self.head_pruning_method:
            weight = self.head_pruning_reshape(weight, self.get_mask(pruning_type="head"))
        return weight, bias

    def forward(self, input, skip_bias_add=False):
        weight, bias = self._effective()
        input = self._quantized_input(input)
        if skip_bias_add:  # tensor-parallel layers add the bias after their all-reduce
            return F.linear(input, weight, None), bias
        return F.linear(input, weight, bias)


class Conv2dLayer_Compress(nn.Conv2d, _CompressState):
    """nn.Conv2d with weight/activation quantisation and sparse / channel pruning (channel = output filter)."""

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self._init_compress_state()

    def __repr__(self):
        s = ("{in_channels}, {out_channels}, kernel_size={kernel_size}, stride={stride}").format(**self.__dict__)
        return "{}({}, sparse pruning={}, channel pruning={}, activation quantization={}, weight_quantization={})".format(
            type(self).__name__, s, self.sparse_pruning_method is not None, self.channel_pruning_method is not None,
            self.activation_quantization_method is not None, self.weight.target_bits)

    def enable_sparse_pruning(self, ratio, method):
        self._enable_pruning("sparse", ratio, method, lambda: self.weight.data.abs(), self.weight.size(), self.weight.size())

    def enable_channel_pruning(self, ratio, method):
        self._enable_pruning("channel", ratio, method, lambda: torch.linalg.norm(self.weight.data, ord=1, dim=[1, 2, 3]),
                             (self.weight.size(0), 1, 1, 1), (-1, 1, 1, 1))

    def fix_channel_pruning_helper(self, mask=None, dim_reduction=False):
        if mask is not None:
            # the conv consuming a pruned conv's output: drop input channels
            self._replace_weight(self.weight.data[:, mask.view(-1), ...])
            self.in_channels = self.weight.size(1)
            self.channel_pruning_enabled = False
            return None
        if self.channel_pruning_method not in ("l1", "topk"):
            raise NotImplementedError
        mask = self.get_mask(pruning_type="channel").bool()
        keep = mask.view(-1)
        if dim_reduction:
            self._replace_weight(self.weight.data[keep, ...])
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data[keep])
            self.out_channels = self.weight.size(0)
        else:
            self.weight.data = self.weight.data * keep.view(-1, 1, 1, 1)
            if self.bias is not None:
                self.bias.data = self.bias.data * keep
        self._drop_pruning_state("channel")
        return mask

    def forward(self, input):
        weight, bias = self._quantized_weight(), self.bias
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            weight = weight * self.get_mask(pruning_type="sparse").view(self.weight.size())
        if self.channel_pruning_enabled:
            mask = self.get_mask(pruning_type="channel")
            weight = weight * mask.view(-1, 1, 1, 1)
            if bias is not None:
                bias = bias * mask.view(-1)
        return F.conv2d(self._quantized_input(input), weight, bias, self.stride, self.padding, self.dilation, self.groups)


class BNLayer_Compress(nn.BatchNorm2d):
    """BatchNorm following a channel-pruned conv: only needs to shrink with it."""

    def fix_channel_pruning_helper(self, mask, dim_reduction=True):
        keep = mask.view(-1)
        self.weight = nn.Parameter(self.weight.data[keep])
        self.bias = nn.Parameter(self.bias.data[keep])
        self.running_mean = self.running_mean[keep]
        self.running_var = self.running_var[keep]
        self.num_features = int(keep.sum())


class Embedding_Compress(nn.Embedding):
    """nn.Embedding with (token-wise grouped) weight quantisation."""

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.weight.start_bits = None
        self.weight.target_bits = None
        self.weight.q_period = None
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False

    def extra_repr(self):
        return "num_embeddings={}, embedding_dim={}, weight_quantization={}".format(self.num_embeddings, self.embedding_dim, self.weight.target_bits)

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits = start_bits
        self.weight.target_bits = target_bits
        self.weight.q_period = quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if weight_quantization_enabled_in_forward:
            logger.warning("************ A lot of MoQ features are not supported in quantize_weight_in_forward mode, please consider to use DS-FP16 optimizer************")
            self.weight_quantizer = _pick_weight_quantizer(target_bits, quantization_type)
            # one group per token row, as upstream forces for embeddings
            self.weight_quantize_num_groups = self.weight.size(0)

    def fix_weight_quantization(self):
        self.weight.data = self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False
        return None

    def forward(self, input):
        weight = self.weight
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            weight = self.weight_quantizer(self.weight, self.weight.target_bits, None, None, self.weight_quantize_num_groups)
        return F.embedding(input, weight, self.padding_idx, self.max_norm, self.norm_type, self.scale_grad_by_freq, self.sparse)


# ---- tensor-parallel (Megatron mpu) variants ------------------------------------------------------------------------
def _mp_group():
    return g_mpu.get_model_parallel_group()


class _MPRegion(torch.autograd.Function):
    """forward / backward chosen from {identity, all-reduce, split-last-dim, gather-last-dim}."""

    @staticmethod
    def forward(ctx, x, fwd, bwd):
        ctx.bwd = bwd
        return _MP_OPS[fwd](x)

    @staticmethod
    def backward(ctx, g):
        return _MP_OPS[ctx.bwd](g), None, None


def _reduce(x):
    if dist.get_world_size(group=_mp_group()) > 1:
        dist.all_reduce(x, group=_mp_group())
    return x


def split_tensor_along_last_dim(tensor, num_partitions, contiguous_split_chunks=False):
    assert tensor.size(-1) % num_partitions == 0
    chunks = torch.split(tensor, tensor.size(-1) // num_partitions, dim=-1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks


def _split(x):
    world = dist.get_world_size(group=_mp_group())
    if world == 1:
        return x
    return split_tensor_along_last_dim(x, world)[dist.get_rank(group=_mp_group())].contiguous()


def _gather(x):
    world = dist.get_world_size(group=_mp_group())
    if world == 1:
        return x
    parts = [torch.empty_like(x) for _ in range(world)]
    parts[dist.get_rank(group=_mp_group())] = x
    dist.all_gather(parts, x, group=_mp_group())
    return torch.cat(parts, dim=-1).contiguous()


_MP_OPS = {"id": lambda t: t, "reduce": _reduce, "split": _split, "gather": _gather}


def copy_to_model_parallel_region(input_):
    return _MPRegion.apply(input_, "id", "reduce")


def reduce_from_model_parallel_region(input_):
    return _MPRegion.apply(input_, "reduce", "id")


def scatter_to_model_parallel_region(input_):
    return _MPRegion.apply(input_, "split", "gather")


def gather_from_model_parallel_region(input_):
    return _MPRegion.apply(input_, "gather", "split")


class ColumnParallelLinear_Compress(LinearLayer_Compress):

    def __init__(self, mpu, input_size, output_size, bias=True, gather_output=True, skip_bias_add=False):
        global g_mpu
        g_mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.gather_output, self.skip_bias_add = gather_output, skip_bias_add
        world_size = mpu.get_model_parallel_world_size()
        assert output_size % world_size == 0
        self.output_size_per_partition = output_size // world_size
        super().__init__(self.input_size, self.output_size_per_partition, bias=bias)

    def forward(self, input_):
        output_parallel, bias = super().forward(copy_to_model_parallel_region(input_), True)
        if bias is not None and not self.skip_bias_add:
            output_parallel = output_parallel + bias
        output = gather_from_model_parallel_region(output_parallel) if self.gather_output else output_parallel
        return output, (bias if self.skip_bias_add else None)


class RowParallelLinear_Compress(LinearLayer_Compress):

    def __init__(self, mpu, input_size, output_size, bias=True, input_is_parallel=False, skip_bias_add=False):
        global g_mpu
        g_mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel, self.skip_bias_add = input_is_parallel, skip_bias_add
        world_size = mpu.get_model_parallel_world_size()
        assert input_size % world_size == 0
        self.input_size_per_partition = input_size // world_size
        super().__init__(self.input_size_per_partition, self.output_size, bias=bias)

    def forward(self, input_):
        input_parallel = input_ if self.input_is_parallel else scatter_to_model_parallel_region(input_)
        output_parallel, bias = super().forward(input_parallel, True)
        output = reduce_from_model_parallel_region(output_parallel)
        if self.skip_bias_add:
            return output, bias
        return (output + bias if bias is not None else output), None
