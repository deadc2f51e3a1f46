"""TiledLinear: split a huge nn.Linear into a grid of smaller ones so ZeRO-3 only materialises one tile at a time.
Parity: deepspeed/runtime/zero/tiling.py:32-296."""
import torch


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    last_dim = tensor.dim() - 1
    tensor_list = torch.split(tensor, partitions, dim=last_dim)
    if contiguous_split_chunks:
        return tuple(chunk.contiguous() for chunk in tensor_list)
    return tensor_list


def _partition_uniform(n, parts):
    base, rem = divmod(n, parts)
    out = [0]
    for i in range(parts):
        out.append(out[-1] + base + (1 if i < rem else 0))
    return out


class TiledLinear(torch.nn.Module):

    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=torch.nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if (in_splits < 1) or (in_splits > in_features):
            raise RuntimeError("in splits must be in range [1, in_features].")
        if (out_splits < 1) or (out_splits > out_features):
            raise RuntimeError("out splits must be in range [1, out_features].")
        self.in_features, self.out_features, self.use_bias = in_features, out_features, bias
        self.out_splits, self.in_splits = out_splits, in_splits
        self.input_is_already_split = input_is_already_split
        self.combine_out_splits = combine_out_splits
        self.in_parts = _partition_uniform(in_features, in_splits)
        self.out_parts = _partition_uniform(out_features, out_splits)
        self.linears = torch.nn.ModuleList()
        for out_id in range(out_splits):
            self.linears.append(torch.nn.ModuleList())
            local_out_dim = self.out_parts[out_id + 1] - self.out_parts[out_id]
            for in_id in range(in_splits):
                local_bias = bias if in_id == (in_splits - 1) else False
                local_in_dim = self.in_parts[in_id + 1] - self.in_parts[in_id]
                self.linears[out_id].append(linear_cls(local_in_dim, local_out_dim, bias=local_bias, **kwargs))
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def forward(self, input_):
        if self.in_splits > 1 and not self.input_is_already_split:
            split_sizes = [self.in_parts[p + 1] - self.in_parts[p] for p in range(self.in_splits)]
            inputs = self._split_global_input(input_, split_sizes)
        elif self.in_splits > 1:
            inputs = input_
            assert len(inputs) == self.in_splits, f"Col splits {self.in_splits} does not match input splits {len(inputs)}"
        else:
            inputs = [input_]
        outputs = [None] * self.out_splits
        for out_id in range(self.out_splits):
            for in_id in range(self.in_splits):
                local_output = self.linears[out_id][in_id](inputs[in_id])
                outputs[out_id] = self._reduce_local_output(in_id, out_id, outputs[out_id], local_output)
        if self.combine_out_splits:
            return self._combine_output_splits(outputs)
        return outputs

    def _split_global_input(self, input, split_sizes):
        return split_tensor_along_last_dim(input, split_sizes)

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        return new_out if current_out is None else current_out + new_out

    def _combine_output_splits(self, outputs):
        assert len(outputs) == self.out_splits
        return torch.cat(outputs, dim=-1)

    @torch.no_grad()
    def copy_params_from(self, other):
        assert hasattr(other, "weight")
        assert other.weight.size() == (self.out_features, self.in_features)
        for row in range(self.out_splits):
            rstart, rstop = self.out_parts[row], self.out_parts[row + 1]
            for col in range(self.in_splits):
                cstart, cstop = self.in_parts[col], self.in_parts[col + 1]
                local = self.linears[row][col]
                local.weight.copy_(other.weight[rstart:rstop, cstart:cstop])
                if local.bias is not None:
                    local.bias.copy_(other.bias[rstart:rstop])


class TiledLinearReturnBias(TiledLinear):
    """Variant for Megatron-style linears that return (output, bias)."""

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        if current_out is None:
            return new_out
        old_tensor, old_bias = current_out
        new_tensor, new_bias = new_out
        assert old_tensor is not None and new_tensor is not None
        tensor = old_tensor + new_tensor
        if old_bias is None:
            bias = new_bias
        elif new_bias is None:
            bias = old_bias
        else:
            bias = old_bias + new_bias
        return tensor, bias

    def _combine_output_splits(self, outputs):
        tensors = [o[0] for o in outputs]
        tensor = super()._combine_output_splits(tensors)
        biases = [o[1] for o in outputs if o[1] is not None]
        bias = super(TiledLinearReturnBias, self)._combine_output_splits(biases) if biases else None
        return tensor, bias
