"""AutoTP: automatic tensor parallelism for arbitrary HF-style transformer modules.
Parity: deepspeed/module_inject/auto_tp.py:187-483 -- `tp_parser` walks the ModuleList blocks and finds the linears
whose outputs must be all-reduced (the last linear before a norm / residual: o_proj, out_proj, down_proj, dense_4h_to_h,
...); `_replace` row-slices those into `LinearAllreduce`, column-slices every other linear into `LinearLayer`
(fused-QKV aware), slices embeddings and updates head-count attributes."""
import re

import torch
from torch import nn

from .. import comm as dist
from ..accelerator import get_accelerator
from .fusedqkv_utils import prepare_tp_fused_qkvw, require_tp_fused_qkvw
from .layers import LinearAllreduce, LinearLayer, LmHeadLinearAllreduce


class ReplaceWithTensorSlicing:
    """Copies a full (un-sharded) tensor into a destination parameter that may be this rank's tensor-parallel slice of it
    (parity: deepspeed/module_inject/auto_tp.py:22-119).  Which dimension is cut follows from comparing the shapes; weights are
    in nn.Linear layout [out, in]."""

    def __init__(self, mp_group=None, mp_size=1, out_dim=1, in_dim=0):
        self.gpu_index = dist.get_rank(group=mp_group) if mp_group is not None else 0
        self.out_dim = out_dim
        self.in_dim = in_dim
        self.mp_size = mp_size

    def merge_assert(self, dim1, dim2):
        assert dim1 > dim2, "Merging tensors is not allowed here! Please use deepspeed load_checkpoint for merging your checkpoints " \
            "before replacing the transformer layer with inference-kernels"

    def _param(self, t):
        return torch.nn.parameter.Parameter(t, requires_grad=False)

    def strided_copy(self, dst, src, num_splits: int, int8=False, allocate_tensor=False):
        """`src` is `num_splits` matrices stacked along dim 0 (fused QKV: 3, [gate; up]: 2): every part is cut alike, so the
        slices of one rank stay aligned (its q heads with its k and v heads)."""
        if src is None:
            return src
        if dst is not None and tuple(src.shape) == tuple(dst.shape):
            return self._store(dst, src)
        parts = torch.chunk(src, num_splits, dim=0)
        if dst is not None and src.dim() > 1 and src.shape[0] == dst.shape[0]:  # row-parallel use of a fused tensor: cut the input dim
            return self._store(dst, torch.cat([p.chunk(self.mp_size, dim=1)[self.gpu_index] for p in parts], dim=0))
        for p in parts:
            assert p.shape[0] % self.mp_size == 0, f"fused tensor part of {p.shape[0]} rows is not divisible by mp_size {self.mp_size}"
        return self._store(dst, torch.cat([p.chunk(self.mp_size, dim=0)[self.gpu_index] for p in parts], dim=0))

    qkv_copy = strided_copy

    def copy(self, dst, src, int8=False, allocate_tensor=False):
        if src is None:
            return src
        if dst is None or allocate_tensor:  # no destination to compare with: column-parallel slice
            return self._param(src.chunk(self.mp_size, dim=0)[self.gpu_index].clone())
        if tuple(src.shape) == tuple(dst.shape):
            return self._store(dst, src)
        if src.dim() == 1:
            self.merge_assert(src.shape[0], dst.shape[0])
            return self._store(dst, src.chunk(self.mp_size, dim=0)[self.gpu_index])
        if src.shape[0] != dst.shape[0]:  # output features differ: column-parallel
            self.merge_assert(src.shape[0], dst.shape[0])
            return self._store(dst, src.chunk(self.mp_size, dim=0)[self.gpu_index])
        self.merge_assert(src.shape[1], dst.shape[1])  # input features differ: row-parallel
        return self._store(dst, src.chunk(self.mp_size, dim=1)[self.gpu_index])

    def _store(self, dst, piece):
        if dst is None:
            return self._param(piece.clone())
        assert tuple(dst.shape) == tuple(piece.shape), f"slice {tuple(piece.shape)} does not fit destination {tuple(dst.shape)}"
        with torch.no_grad():
            dst.data.copy_(piece.to(dst.dtype))
        return dst if isinstance(dst, torch.nn.Parameter) else self._param(dst)


class Loading:

    @staticmethod
    def is_load_module(module):
        load_layers = [nn.Linear, nn.Embedding, nn.LayerNorm]
        load_layer_names = ["LPLayerNorm", "SharedEmbedding", "OPTLearnedPositionalEmbedding", "LlamaRMSNorm", "FalconLinear",
                            "MistralRMSNorm", "T5LayerNorm", "MixtralRMSNorm"]
        return module.__class__ in load_layers or module._get_name() in load_layer_names


class AutoTP:

    def __init__(self, module, all_reduce_linears, prefix, state_dict, linear_layer_setting, orig_layer_impl):
        self.module = module
        self.all_reduce_linears = all_reduce_linears
        self.prefix = prefix
        self.state_dict = state_dict
        self.mp_size = None
        self.mp_group = None
        self.linear_layer_setting = linear_layer_setting
        self.orig_layer_impl = orig_layer_impl
        self.linear_policies = None
        self.conv_linear_layer = False

    # ------------------------------------------------------------------ parsing: which linears end in an all-reduce
    # A transformer block is flattened into the ordered token list of its leaves: ".attn.o_proj" for a Linear (qualified by
    # the names of the two innermost containers, as the policies expect), "ln" for a normalisation layer.  A linear needs
    # an all-reduce when (a) it is the last linear before a norm, (b) its name is a known row-parallel output projection, or
    # (c) it is the last linear of an MLP-like child that matched nothing else.
    _NORM_NAMES = ("LayerNorm", "layer_norm", "ln_1", "ln_2")
    _UNSUPPORTED = ("deberta", "flaubert", "fsmt", "gpt2", "led", "longformer", "xlm", "xlnet")

    @staticmethod
    def in_module_list(module, module_list):
        return any(type(m).__name__ == type(module).__name__ for m in module_list)

    @staticmethod
    def get_module_list(model):
        """One representative of every distinct block class found inside the model's nn.ModuleLists (depth first)."""
        found = []
        for child in model.children():
            candidates = child.children() if isinstance(child, nn.ModuleList) else AutoTP.get_module_list(child)
            for block in candidates:
                if not AutoTP.in_module_list(block, found):
                    found.append(block)
        return found

    @staticmethod
    def supported(model):
        text = str(model)
        for pattern, fn in ((r": (.*?)Model", re.search), (r": (.*?)Stack", re.search), (r"(.*?)Model", re.match)):
            hit = fn(pattern, text)
            if hit is not None:
                return hit.group(1).lower() not in AutoTP._UNSUPPORTED
        return True

    @staticmethod
    def _is_norm(key, module):
        return isinstance(module, nn.LayerNorm) or key in AutoTP._NORM_NAMES[:2]

    @staticmethod
    def get_layers(parent, module):
        tokens = []
        for key, sub in module._modules.items():
            if isinstance(sub, nn.Linear):
                tokens.append(f"{parent}.{key}")
            elif AutoTP._is_norm(key, sub):
                tokens.append("ln")
            else:
                tokens.extend(AutoTP.get_layers(key, sub))
        return tokens

    @staticmethod
    def update_policy_list(policy_list, new_module, new_gems):
        for i, (cls, gems) in enumerate(policy_list):
            if cls is type(new_module):
                policy_list[i] = (cls, set(new_gems) | set(gems))
                return policy_list
        policy_list.append((type(new_module), new_gems))
        return policy_list

    @staticmethod
    def kernel_supported(module_list):
        return False

    @staticmethod
    def _is_output_projection(token, model_text, block_type_text):
        if any(tag in token for tag in ("out_proj", "o_proj", "down_proj", "dense_4h_to_h")):
            return True
        if "attention.dense" in token and "GPTNeoX" in model_text:
            return True
        if "self_attention.dense" in token and "falcon" in block_type_text:
            return True
        return "w2" in token and "mixtral" in block_type_text

    @staticmethod
    def tp_parser(model):
        """[(block class, names of the linears whose output must be all-reduced)]"""
        unsupported_msg = "AutoTP not supported for model. Please use kernel injection since container policy for model exists."
        assert AutoTP.supported(model), unsupported_msg
        model_text = str(model)
        policies = []
        for block in AutoTP.get_module_list(model):
            block_text = str(type(block)).lower()
            for key, sub in block._modules.items():
                if isinstance(sub, nn.Linear):
                    tokens = ["." + key]
                elif isinstance(sub, nn.LayerNorm) or key in AutoTP._NORM_NAMES:
                    tokens = ["ln"]
                else:
                    tokens = AutoTP.get_layers(key, sub)
                picked = set()
                for before, tok in zip([None] + tokens[:-1], tokens):
                    if tok == "ln":
                        if before not in (None, "ln"):
                            picked.add(before)
                    elif AutoTP._is_output_projection(tok, model_text, block_text):
                        picked.add(tok)
                if not picked and tokens and tokens[-1] != "ln" and key in ("mlp", "feed_forward", "ffn"):
                    picked.add(tokens[-1])
                if picked:
                    policies = AutoTP.update_policy_list(policies, block, list(picked))
        assert policies, unsupported_msg
        return policies

    def set_tensor_parallel_config(self, mp_size, mp_group):
        self.mp_size = mp_size
        self.mp_group = mp_group

    def _rank(self):
        return dist.get_rank(self.mp_group) if (self.mp_group is not None and dist.is_initialized()) else 0

    def _replace(self, child, name, conv_linear_layer=False):
        if getattr(child, "replaced", False):
            return child
        weight_shape = child.weight.shape
        mp, r = self.mp_size, self._rank()
        dev = get_accelerator().current_device_name()
        if name in self.all_reduce_linears:
            # row parallel: slice input features, all-reduce output
            data = child.weight.data.chunk(mp, dim=1)[r].contiguous().to(dev)
            bias = None if child.bias is None else torch.nn.parameter.Parameter(child.bias.data.to(dev), requires_grad=False)
            new = LinearAllreduce(torch.nn.parameter.Parameter(data, requires_grad=False), bias, self.mp_group)
            new.replaced = True
            return new
        # column parallel
        module_str = str(self.module).strip()
        if require_tp_fused_qkvw(name, mp):
            cfg = getattr(self.module, "config", None)
            data = prepare_tp_fused_qkvw(module_str, child.weight.data, mp, r)
            bias = prepare_tp_fused_qkvw(module_str, child.bias.data, mp, r) if child.bias is not None else None
        else:
            data = child.weight.data.chunk(mp, dim=0)[r]
            bias = child.bias.data.chunk(mp, dim=0)[r] if child.bias is not None else None
        new = LinearLayer(weight=torch.nn.parameter.Parameter(data.contiguous().to(dev), requires_grad=False),
                          bias=None if bias is None else torch.nn.parameter.Parameter(bias.contiguous().to(dev), requires_grad=False))
        new.replaced = True
        return new

    def _slice_embedding(self, child, name, conv_linear_layer=False):
        if getattr(child, "replaced", False):
            return child
        mp, r = self.mp_size, self._rank()
        data = child.weight.data.chunk(mp, dim=1)[r].contiguous().to(get_accelerator().current_device_name())
        new_embedding = nn.Embedding(child.weight.shape[0], data.shape[1])
        new_embedding.weight = torch.nn.parameter.Parameter(data, requires_grad=False)
        new_embedding.replaced = True
        return new_embedding

    _HEAD_COUNT_ATTRS = ("n_heads", "inner_dim", "num_heads", "num_kv", "num_attention_heads", "num_attn_heads", "all_head_size",
                         "embed_dim", "hidden_size", "num_key_value_heads", "num_kv_heads", "kv_n_heads", "d_model")

    def update_mp_params(self, child):
        """Attention modules cache head counts / widths: divide the ones that split evenly, once per module."""
        if getattr(child, "replaced", False):
            return
        for attr in self._HEAD_COUNT_ATTRS:
            value = getattr(child, attr, None)
            if isinstance(value, int) and value % self.mp_size == 0:
                setattr(child, attr, value // self.mp_size)
        child.replaced = True

    def update_linear_policies(self):
        self.conv_linear_layer = False
        setting = self.linear_layer_setting
        if setting is None:
            self.linear_policies = {nn.Linear: self._replace}
            return
        self.linear_policies = {setting[0]: self._replace}
        if len(setting) == 2:
            self.linear_policies[setting[1]] = self._slice_embedding

    def _replace_module(self, r_module, prev_name="", prev_class_name=""):
        """Depth-first walk: leaves with a policy are swapped for their sliced twins, every other module gets its head counts
        adjusted.  A leaf is known to the policies by ".<parent>.<name>"."""
        scope = ".".join(x for x in (prev_class_name, prev_name) if x)
        for name, child in r_module.named_children():
            swap = self.linear_policies.get(type(child))
            if swap is not None:
                setattr(r_module, name, swap(child, f"{prev_name}.{name}", self.conv_linear_layer))
            else:
                self.update_mp_params(child)
                self._replace_module(child, name, scope)
        return r_module

    def get_model_num_kv_heads(self, config):
        num_kv_heads = None
        for name in ["num_kv_heads", "num_key_value_heads", "num_attention_heads", "n_heads"]:
            if hasattr(config, name):
                num_kv_heads = getattr(config, name)
                if num_kv_heads is not None:
                    break
        return num_kv_heads

    def _replace_last_linear_module(self, r_module):
        if hasattr(r_module, "lm_head"):
            name = "lm_head"
            child = r_module.lm_head
        elif hasattr(r_module, "embed_out"):
            name = "embed_out"
            child = r_module.embed_out
        else:
            return r_module
        if isinstance(child, nn.Linear) and child.weight.shape[1] % self.mp_size == 0 and self.mp_size > 1:
            mp, r = self.mp_size, self._rank()
            dev = get_accelerator().current_device_name()
            data = child.weight.data.chunk(mp, dim=1)[r].contiguous().to(dev)
            bias = None if child.bias is None else torch.nn.parameter.Parameter(child.bias.data.to(dev), requires_grad=False)
            setattr(r_module, name, LmHeadLinearAllreduce(torch.nn.parameter.Parameter(data, requires_grad=False), r, mp, bias, self.mp_group))
        return r_module
