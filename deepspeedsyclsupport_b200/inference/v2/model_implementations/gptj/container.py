"""GPT-J rotates interleaved pairs (2i, 2i+1) of the first rotary_dim features of every head; the attention kernels rotate
(i, i + rot/2).  Permuting those rows of W_q and W_k the same way (evens first, then odds) leaves q.k unchanged."""
from typing import Optional

import torch

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, MLP1Parameter, MLP2Parameter, NormParameter,
                                 UnembedParameter, UnfusedQKVParameter)
from ..layer_container_base import LayerContainer


class GPTJTransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    o_w: AttentionOutputParameter
    mlp1_w: MLP1Parameter
    mlp1_b: MLP1Parameter
    mlp2_w: MLP2Parameter
    mlp2_b: MLP2Parameter
    norm1_w: NormParameter
    norm1_b: NormParameter

    PARAM_MAPPING = {
        "attn.q_proj.weight": ("qkv_w.q_params", "to_rotate_half"),
        "attn.k_proj.weight": ("qkv_w.k_params", "to_rotate_half"),
        "attn.v_proj.weight": "qkv_w.v_params",
        "attn.out_proj.weight": "o_w.params",
        "mlp.fc_in.weight": "mlp1_w.params",
        "mlp.fc_in.bias": "mlp1_b.params",
        "mlp.fc_out.weight": "mlp2_w.params",
        "mlp.fc_out.bias": "mlp2_b.params",
        "ln_1.weight": "norm1_w.params",
        "ln_1.bias": "norm1_b.params",
    }

    def to_rotate_half(self, w):
        s = self.inference_model.spec
        d, rot = s.head_dim, s.rotary_dim
        perm = torch.cat([torch.arange(0, rot, 2), torch.arange(1, rot, 2), torch.arange(rot, d)])
        rows = (torch.arange(w.shape[0] // d)[:, None] * d + perm[None, :]).reshape(-1)
        return w[rows]


class GPTJNonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    lm_head: UnembedParameter
    lm_head_b: Optional[UnembedParameter]
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "transformer.wte.weight": "embed.params",
        "transformer.ln_f.weight": "final_norm_w.params",
        "transformer.ln_f.bias": "final_norm_b.params",
        "lm_head.weight": "lm_head.params",
        "lm_head.bias": "lm_head_b.params",
    }
