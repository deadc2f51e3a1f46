"""One ragged-batch decoder implementation parameterised by an `ArchSpec`, covering the model families the reference
ships as separate containers/policies (inference/v2/model_implementations/{llama_v2,mistral,mixtral,opt,falcon,phi}):

  llama / mistral / internlm (attention biases) : RMSNorm, rotary, gated SiLU MLP, GQA           mixtral : + top-2 MoE MLP (experts TP-sharded)
  opt             : LayerNorm, learned positions (+2), ReLU MLP, biases      falcon / phi : parallel attention + MLP block

Forward of a layer: fused (residual-add +) norm -> QKV GEMM -> RoPE + paged-KV append -> paged attention -> output GEMM ->
tensor-parallel all-reduce -> norm -> MLP GEMMs -> all-reduce (reference: llama_v2/model.py:151-213).
Weights are sharded per tensor-parallel rank when they are loaded (column slices of qkv/up/gate, row slices of o/down,
vocab-parallel unembedding) -- reference: model_implementations/sharding/*.py."""
import math
from dataclasses import dataclass, field
from typing import Optional

import torch

from .... import comm as dist
from ....ops import fused_ops as FO
from ....ops.gemm import gemm, moe_gemm
from .. import kernels as K


@dataclass
class ArchSpec:
    family: str = "llama"
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_layers: int = 32
    num_heads: int = 32
    num_kv_heads: int = 32
    head_dim: int = 128
    norm: str = "rms"            # rms | ln
    norm_eps: float = 1e-5
    activation: str = "silu"     # gated activations: silu/gelu/relu (gated=True) ; plain: gelu/relu (gated=False)
    gated_mlp: bool = True
    rotary_dim: int = 128        # 0 = no rotary
    rope_theta: float = 10000.0
    learned_pos_offset: Optional[int] = None   # OPT: 2
    parallel_block: bool = False               # falcon / phi / gpt-j: attention and MLP read the same normed input
    parallel_norm2: bool = False               # gpt-neox: parallel residual, but the MLP has its own LayerNorm of the input
    attn_bias: bool = False
    mlp_bias: bool = False
    num_experts: int = 0
    top_k: int = 2
    tie_embeddings: bool = False
    max_positions: int = 4096


def _rope(c, default_partial=1.0):
    """(theta, partial rotary factor) from either the flat keys of older HF configs or the `rope_parameters` dict of newer ones."""
    rp = c.get("rope_parameters") or {}
    theta = c.get("rope_theta", c.get("rotary_emb_base", rp.get("rope_theta", 10000.0)))
    partial = c.get("partial_rotary_factor", c.get("rotary_pct", rp.get("partial_rotary_factor", default_partial)))
    return float(theta), float(partial)


def _spec_llama(c):
    return ArchSpec(family="llama", vocab_size=c["vocab_size"], hidden_size=c["hidden_size"], intermediate_size=c["intermediate_size"],
                    num_layers=c["num_hidden_layers"], num_heads=c["num_attention_heads"],
                    num_kv_heads=c.get("num_key_value_heads", c["num_attention_heads"]),
                    head_dim=c["hidden_size"] // c["num_attention_heads"], norm_eps=c.get("rms_norm_eps", 1e-5),
                    rotary_dim=c["hidden_size"] // c["num_attention_heads"], rope_theta=_rope(c)[0],
                    tie_embeddings=c.get("tie_word_embeddings", False), max_positions=c.get("max_position_embeddings", 4096))


def _spec_mixtral(c):
    s = _spec_llama(c)
    s.family, s.num_experts, s.top_k = "mixtral", c.get("num_local_experts", 8), c.get("num_experts_per_tok", 2)
    return s


def _spec_opt(c):
    return ArchSpec(family="opt", vocab_size=c["vocab_size"], hidden_size=c["hidden_size"], intermediate_size=c["ffn_dim"],
                    num_layers=c["num_hidden_layers"], num_heads=c["num_attention_heads"], num_kv_heads=c["num_attention_heads"],
                    head_dim=c["hidden_size"] // c["num_attention_heads"], norm="ln", activation="relu", gated_mlp=False, rotary_dim=0,
                    learned_pos_offset=2, attn_bias=True, mlp_bias=True, tie_embeddings=True, max_positions=c.get("max_position_embeddings", 2048))


def _spec_falcon(c):
    hq = c["num_attention_heads"]
    return ArchSpec(family="falcon", vocab_size=c["vocab_size"], hidden_size=c["hidden_size"], intermediate_size=4 * c["hidden_size"],
                    num_layers=c["num_hidden_layers"], num_heads=hq, num_kv_heads=c.get("num_kv_heads", 1 if c.get("multi_query", True) else hq),
                    head_dim=c["hidden_size"] // hq, norm="ln", activation="gelu", gated_mlp=False, rotary_dim=c["hidden_size"] // hq,
                    rope_theta=_rope(c)[0], parallel_block=c.get("parallel_attn", True), attn_bias=c.get("bias", False), mlp_bias=c.get("bias", False))


def _spec_phi(c):
    hq = c["num_attention_heads"]
    d = c["hidden_size"] // hq
    return ArchSpec(family="phi", vocab_size=c["vocab_size"], hidden_size=c["hidden_size"], intermediate_size=c.get("intermediate_size", 4 * c["hidden_size"]),
                    num_layers=c["num_hidden_layers"], num_heads=hq, num_kv_heads=c.get("num_key_value_heads", hq) or hq, head_dim=d, norm="ln",
                    norm_eps=c.get("layer_norm_eps", 1e-5), activation="gelu", gated_mlp=False,
                    rotary_dim=int(d * _rope(c, 0.5)[1]), rope_theta=_rope(c)[0], parallel_block=True, attn_bias=True, mlp_bias=True)


def _spec_gpt2(c):
    h = c["n_embd"]
    return ArchSpec(family="gpt2", vocab_size=c["vocab_size"], hidden_size=h, intermediate_size=c.get("n_inner") or 4 * h,
                    num_layers=c["n_layer"], num_heads=c["n_head"], num_kv_heads=c["n_head"], head_dim=h // c["n_head"], norm="ln",
                    norm_eps=c.get("layer_norm_epsilon", 1e-5), activation="gelu", gated_mlp=False, rotary_dim=0, learned_pos_offset=0,
                    attn_bias=True, mlp_bias=True, tie_embeddings=True, max_positions=c.get("n_positions", 1024))


def _spec_gptj(c):
    h, hq = c["n_embd"], c["n_head"]
    return ArchSpec(family="gptj", vocab_size=c["vocab_size"], hidden_size=h, intermediate_size=c.get("n_inner") or 4 * h,
                    num_layers=c["n_layer"], num_heads=hq, num_kv_heads=hq, head_dim=h // hq, norm="ln",
                    norm_eps=c.get("layer_norm_epsilon", 1e-5), activation="gelu", gated_mlp=False,
                    rotary_dim=c.get("rotary_dim") or h // hq, parallel_block=True, mlp_bias=True, max_positions=c.get("n_positions", 2048))


def _spec_gpt_neox(c):
    h, hq = c["hidden_size"], c["num_attention_heads"]
    d = h // hq
    par = c.get("use_parallel_residual", True)
    return ArchSpec(family="gpt_neox", vocab_size=c["vocab_size"], hidden_size=h, intermediate_size=c.get("intermediate_size", 4 * h),
                    num_layers=c["num_hidden_layers"], num_heads=hq, num_kv_heads=hq, head_dim=d, norm="ln",
                    norm_eps=c.get("layer_norm_eps", 1e-5), activation="gelu", gated_mlp=False,
                    rotary_dim=int(d * _rope(c, 0.25)[1]), rope_theta=_rope(c)[0], parallel_block=par, parallel_norm2=par,
                    attn_bias=True, mlp_bias=True, max_positions=c.get("max_position_embeddings", 2048))


SPEC_FROM_HF = {"llama": _spec_llama, "mistral": _spec_llama, "internlm": _spec_llama, "mixtral": _spec_mixtral, "opt": _spec_opt,
                "falcon": _spec_falcon, "phi": _spec_phi, "gpt2": _spec_gpt2, "gptj": _spec_gptj, "gpt_neox": _spec_gpt_neox}


def _shard_rows(w, rank, world):
    """column-parallel linear: slice output features"""
    n = w.shape[0] // world
    return w[rank * n:(rank + 1) * n].contiguous()


def _shard_cols(w, rank, world):
    """row-parallel linear: slice input features"""
    n = w.shape[1] // world
    return w[:, rank * n:(rank + 1) * n].contiguous()


class RaggedTransformer:
    """Holds the TP shard of the weights and runs ragged batches against a DSStateManager."""

    def __init__(self, spec: ArchSpec, state_dict, tp_group=None, dtype=torch.bfloat16, device=None, symm=None, quantization_mode=""):
        self.spec = spec
        self.quantization_mode = quantization_mode or ""
        self.tp_group = tp_group
        self.tp = dist.get_world_size(tp_group) if (tp_group is not None and dist.is_initialized()) else 1
        self.rank = dist.get_rank(tp_group) if self.tp > 1 else 0
        self.dtype = dtype
        self.device = device or torch.device("cpu")
        self.symm = symm
        s = spec
        assert s.num_heads % self.tp == 0, "num_heads must be divisible by tp_size"
        self.hq = s.num_heads // self.tp
        self.hkv = max(1, s.num_kv_heads // self.tp)
        self.kv_replicated = s.num_kv_heads < self.tp
        self.inv_freq = FO.rope_inv_freq(s.rotary_dim, s.rope_theta, device=self.device) if s.rotary_dim else None
        self._load(state_dict)

    # ---- weight loading (canonical names; engine_factory maps HF names onto these)
    def _t(self, x):
        return x.to(self.device, self.dtype).contiguous()

    def _load(self, sd):
        s, r, W = self.spec, self.rank, self.tp
        d = s.head_dim
        self.embed = self._t(sd["embed.weight"])
        self.pos_embed = self._t(sd["pos_embed.weight"]) if "pos_embed.weight" in sd else None
        self.final_norm_w = self._t(sd["final_norm.weight"])
        self.final_norm_b = self._t(sd["final_norm.bias"]) if "final_norm.bias" in sd else None
        head = sd["embed.weight"] if s.tie_embeddings or "lm_head.weight" not in sd else sd["lm_head.weight"]
        vpad = -(-head.shape[0] // W) * W
        if vpad != head.shape[0]:
            head = torch.cat([head, head.new_zeros(vpad - head.shape[0], head.shape[1])])
        self.lm_head = self._t(_shard_rows(head, r, W))
        self.lm_head_b = None
        if "lm_head.bias" in sd:  # gpt-j
            hb = sd["lm_head.bias"]
            if vpad != hb.shape[0]:
                hb = torch.cat([hb, hb.new_zeros(vpad - hb.shape[0])])
            self.lm_head_b = self._t(hb.view(W, -1)[r])
        self.layers = []
        for i in range(s.num_layers):
            p = f"layers.{i}."
            L = {}
            q, k, v = sd[p + "q.weight"], sd[p + "k.weight"], sd[p + "v.weight"]
            qs = _shard_rows(q, r, W)
            if self.kv_replicated:
                g = r * s.num_kv_heads // W
                ks, vs = k[g * d:(g + 1) * d], v[g * d:(g + 1) * d]
            else:
                ks, vs = _shard_rows(k, r, W), _shard_rows(v, r, W)
            L["qkv_w"] = self._t(torch.cat([qs, ks, vs], 0))
            if p + "q.bias" in sd:
                qb, kb, vb = sd[p + "q.bias"], sd[p + "k.bias"], sd[p + "v.bias"]
                qbs = qb.view(W, -1)[r]
                kbs = kb[g * d:(g + 1) * d] if self.kv_replicated else kb.view(W, -1)[r]
                vbs = vb[g * d:(g + 1) * d] if self.kv_replicated else vb.view(W, -1)[r]
                L["qkv_b"] = self._t(torch.cat([qbs, kbs, vbs]))
            L["o_w"] = self._t(_shard_cols(sd[p + "o.weight"], r, W))
            if p + "o.bias" in sd:
                L["o_b"] = self._t(sd[p + "o.bias"])
            L["norm1_w"] = self._t(sd[p + "norm1.weight"])
            if p + "norm1.bias" in sd:
                L["norm1_b"] = self._t(sd[p + "norm1.bias"])
            if p + "norm2.weight" in sd:
                L["norm2_w"] = self._t(sd[p + "norm2.weight"])
                if p + "norm2.bias" in sd:
                    L["norm2_b"] = self._t(sd[p + "norm2.bias"])
            if s.num_experts:
                L["router_w"] = sd[p + "router.weight"].to(self.device, torch.float32).contiguous()
                # expert weights stacked [E, N, K]: one grouped GEMM launch covers every expert (ops/gemm.moe_gemm)
                gus, dns = [], []
                for e in range(s.num_experts):
                    ep = f"{p}experts.{e}."
                    gus.append(torch.cat([_shard_rows(sd[ep + "gate.weight"], r, W), _shard_rows(sd[ep + "up.weight"], r, W)], 0))
                    dns.append(_shard_cols(sd[ep + "down.weight"], r, W))
                L["experts_gu"] = self._t(torch.stack(gus))
                L["experts_dn"] = self._t(torch.stack(dns))
            elif s.gated_mlp:
                L["mlp1_w"] = self._t(torch.cat([_shard_rows(sd[p + "gate.weight"], r, W), _shard_rows(sd[p + "up.weight"], r, W)], 0))
                L["mlp2_w"] = self._t(_shard_cols(sd[p + "down.weight"], r, W))
            else:
                L["mlp1_w"] = self._t(_shard_rows(sd[p + "up.weight"], r, W))
                if p + "up.bias" in sd:
                    L["mlp1_b"] = self._t(sd[p + "up.bias"].view(W, -1)[r])
                L["mlp2_w"] = self._t(_shard_cols(sd[p + "down.weight"], r, W))
                if p + "down.bias" in sd:
                    L["mlp2_b"] = self._t(sd[p + "down.bias"])
            if self.quantization_mode == "wf8af8":
                # W8A8: dense projection weights stored as e4m3 with one scale per output feature; activations are quantised per
                # token inside _linear and the product runs on the fp8 tensor-core GEMM (csrc/gemm/gemm_fp8.cu)
                from ....ops.gemm import quantize_fp8_rowwise
                for key in ("qkv_w", "o_w", "mlp1_w", "mlp2_w"):
                    if key in L and L[key].shape[1] % 16 == 0:
                        L[key] = quantize_fp8_rowwise(L[key])
            self.layers.append(L)

    # ---- building blocks
    def _linear(self, x, w, b=None):
        if isinstance(w, tuple):  # (e4m3 weight, per-row scale)
            from ....ops.gemm import linear_fp8
            y = linear_fp8(x, w[0], w[1])
            return y if b is None else y + b
        y = gemm(x, w) if x.is_cuda else torch.nn.functional.linear(x, w)
        if b is not None:
            y = y + b
        return y

    def _norm(self, x, w, b, residual=None):
        if self.spec.norm == "rms":
            return FO.rms_norm(x, w, self.spec.norm_eps, residual)
        return FO.layer_norm(x, w, b, self.spec.norm_eps, residual)

    def _all_reduce(self, x):
        if self.tp > 1:
            dist.all_reduce(x, group=self.tp_group)
        return x

    def _mlp(self, L, h):
        s = self.spec
        if s.num_experts:
            logits = torch.nn.functional.linear(h.float(), L["router_w"])
            counts, scores, asg, off = K.top_k_gating(logits, s.top_k, normalize=True)
            moe_in, cumsum, slots = K.moe_scatter(h, counts, asg, off)
            # the row ranges stay on the device: no host synchronisation between gating and the expert GEMMs
            act = FO.gated_act(moe_gemm(moe_in, L["experts_gu"], cumsum), s.activation)
            return K.moe_gather(moe_gemm(act, L["experts_dn"], cumsum), scores, slots)
        if s.gated_mlp:
            return self._linear(FO.gated_act(self._linear(h, L["mlp1_w"]), s.activation), L["mlp2_w"])
        a = FO.bias_act(self._linear(h, L["mlp1_w"]), L.get("mlp1_b"), s.activation)
        return self._linear(a, L["mlp2_w"])

    def forward(self, batch, state_manager):
        """batch: RaggedBatchWrapper (finalized).  Returns logits [n_seqs, vocab] for the last token of every sequence."""
        s = self.spec
        ids, tok_seq, tok_pos, bt = batch.input_ids(), batch.tok_seq(), batch.tok_pos(), batch.block_table()
        x = K.ragged_embed(ids, self.embed, self.pos_embed, tok_pos, s.learned_pos_offset or 0)
        residual = None
        for li, L in enumerate(self.layers):
            if residual is None:
                residual = x
                hn = self._norm(x, L["norm1_w"], L.get("norm1_b"))
            else:
                hn, residual = self._norm(x, L["norm1_w"], L.get("norm1_b"), residual)
            qkv = self._linear(hn, L["qkv_w"], L.get("qkv_b"))
            cache = state_manager.get_cache(li)
            K.kv_rotary_append(qkv, cache, tok_seq, tok_pos, bt, self.inv_freq, self.hq, self.hkv, s.head_dim, rot_dim=s.rotary_dim)
            attn = K.paged_attention(qkv, cache, tok_seq, tok_pos, bt, self.hq, self.hkv, s.head_dim)
            attn_out = self._linear(attn, L["o_w"])
            if s.parallel_block:
                # gpt-neox gives the MLP its own LayerNorm of the block input (= the residual stream after the fused add)
                mlp_in = self._norm(residual, L["norm2_w"], L.get("norm2_b")) if s.parallel_norm2 else hn
                mlp_out = self._mlp(L, mlp_in)
                x = self._all_reduce(attn_out + mlp_out)
                if "o_b" in L:
                    x = x + L["o_b"]
                if "mlp2_b" in L:
                    x = x + L["mlp2_b"]
                continue
            attn_out = self._all_reduce(attn_out)
            if "o_b" in L:
                attn_out = attn_out + L["o_b"]
            hn, residual = self._norm(attn_out, L["norm2_w"], L.get("norm2_b"), residual)
            x = self._all_reduce(self._mlp(L, hn))
            if "mlp2_b" in L:
                x = x + L["mlp2_b"]
        last = K.logits_gather(x, batch.last_tokens())
        last_res = K.logits_gather(residual, batch.last_tokens())
        hn, _ = self._norm(last, self.final_norm_w, self.final_norm_b, last_res)
        logits = self._linear(hn, self.lm_head, self.lm_head_b)
        if self.tp > 1:
            full = torch.empty(self.tp, *logits.shape, dtype=logits.dtype, device=logits.device)
            dist.all_gather_into_tensor(full, logits.contiguous(), group=self.tp_group)
            logits = full.permute(1, 0, 2).reshape(logits.shape[0], -1)
        return logits[:, :s.vocab_size]
