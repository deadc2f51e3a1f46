"""WeightQuantization: symmetric int8 group-wise quantisation of transformer weights, either while a Megatron
checkpoint is merged / split for a new tensor-parallel degree (state_dict_factory) or on a live model through
injection policies.  Besides the int8 tensors it produces, per layer, one row of dequantisation scales laid out as
[qkv | attention-out | mlp-in | mlp-out] (zero padded to equal width), which is the table the quantised inference
kernels index.  API parity: deepspeed/runtime/weight_quantizer.py."""
import torch

_KINDS = ("qkv", "dense", "h4h", "4hh")  # scale-table column order


def _kind_of(key):
    if "mlp.dense_4h_to_h.weight" in key:
        return "4hh"
    if "mlp.dense_h_to_4h.weight" in key:
        return "h4h"
    if "attention.query_key_value.weight" in key:
        return "qkv"
    return "dense"


class WeightQuantization(object):

    def __init__(self, mlp_extra_grouping=True, mp_size=1):
        self._scales = {k: [] for k in _KINDS}
        self.mlp_extra_grouping = mlp_extra_grouping
        self.mp_size = mp_size

    # legacy attribute names
    dense_scales = property(lambda self: self._scales["dense"])
    qkv_scales = property(lambda self: self._scales["qkv"])
    mlp4hh_scales = property(lambda self: self._scales["4hh"])
    mlph4h_scales = property(lambda self: self._scales["h4h"])

    def quantize_data(self, data, quantize_bits, groups, key=None):
        """-> (int8 tensor of data's shape, per-group scale with  q = round(x * scale))."""
        levels = 2**quantize_bits
        g = data.float().reshape(groups, -1)
        scale = levels / (2 * g.abs().amax(dim=1, keepdim=True) + 1e-5)
        q = (g * scale).round().clamp(-levels // 2, levels // 2 - 1)
        return q.reshape(data.shape).to(torch.int8).contiguous(), scale.reshape(-1)

    def _ratio_is(self, data, ratio, merge_count=1):
        rows, cols = data.shape[0], data.shape[1]
        return (self.mp_size * rows * merge_count) / cols == ratio or (self.mp_size * cols * merge_count) / rows == ratio

    def is_mlp(self, data, merge_count=1):
        return self._ratio_is(data, 4, merge_count)

    def is_qkv(self, data):
        return self._ratio_is(data, 3)

    def Quantize(self, value_list, quantize_bits, groups, key, merge_dim=0):
        """Quantise every shard of one parameter in place (list is modified and returned); remember its scales."""
        if self.mlp_extra_grouping and self.is_mlp(value_list[0], merge_count=len(value_list)):
            groups *= 2
        shard_scales = []
        for i, shard in enumerate(value_list):
            value_list[i], s = self.quantize_data(shard, quantize_bits, groups, key)
            shard_scales.append(s)
        inv = 1.0 / torch.cat(shard_scales, dim=0).to(torch.float32).view(1, -1)  # stored as dequantisation factors
        self._scales[_kind_of(key)].append(inv)
        return value_list

    def merge_layer_scales(self, layer_scales):
        width = max(s.shape[-1] for s in layer_scales)
        rows = [torch.nn.functional.pad(s, (0, width - s.shape[-1])) for s in layer_scales]
        return torch.cat(rows).unsqueeze(0)

    def _per_layer(self):
        return zip(*(self._scales[k] for k in _KINDS))

    def merge_scales(self):
        return torch.cat([self.merge_layer_scales(list(layer)) for layer in self._per_layer()])

    def merge_scales_split(self, split_count):
        per_rank = [[] for _ in range(split_count)]
        for qkv, dense, h4h, m4hh in self._per_layer():
            parts = {name: torch.split(t, t.numel() // split_count, dim=-1) for name, t in (("qkv", qkv), ("dense", dense), ("h4h", h4h), ("4hh", m4hh))}
            for r in range(split_count):
                widen = lambda t: torch.cat((t, torch.zeros_like(t)), dim=1)  # qkv / dense rows are half as wide as the mlp rows
                per_rank[r].append(torch.cat([widen(parts["qkv"][r]), widen(parts["dense"][r]), parts["h4h"][r], parts["4hh"][r]]).unsqueeze(0))
        return per_rank

    def sd_quantize_megatron(self, sd, quantize_bits, groups):
        targets = ("attention.dense.weight", "mlp.dense_4h_to_h.weight", "mlp.dense_h_to_4h.weight", "attention.query_key_value.weight")
        for key in sd.keys():
            if any(t in key for t in targets):
                sd[key] = self.Quantize([sd[key]], quantize_bits, groups, key=key)[0]
        return sd, self.merge_scales()

    def model_quantize(self, model, quantize_policy, quantize_bits, groups):
        """Quantise, in place, the four weight matrices every policy-matched layer exposes; returns (model, scales)."""
        all_scales = []

        def quantize_layer(layer, policy_cls):
            policy = policy_cls(layer)
            _, qkvw, _, dense_w, _, _ = policy.attention()
            _, h4h_w, _, m4hh_w, _ = policy.mlp()
            layer_scales = []
            for w in (qkvw, dense_w, h4h_w, m4hh_w):
                g = groups
                if self.mlp_extra_grouping and self.is_mlp(w):
                    g = groups * 2
                elif self.is_qkv(w):
                    g = groups * 3
                q, s = self.quantize_data(w, quantize_bits, g)
                w.copy_(q)
                layer_scales.append(1.0 / s.to(torch.float32).view(1, -1))
            all_scales.append(self.merge_layer_scales(layer_scales))
            return layer

        policies = dict(quantize_policy or {})

        def walk(module):
            for name, child in module.named_children():
                if child.__class__ in policies:
                    setattr(module, name, quantize_layer(child, policies[child.__class__]))
                else:
                    walk(child)
            return module

        return walk(model), torch.cat(all_scales)
