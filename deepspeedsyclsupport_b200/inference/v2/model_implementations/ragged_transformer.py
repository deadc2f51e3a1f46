"""One ragged-batch decoder implementation parameterised by an `ArchSpec`, covering the model families the reference
ships as separate containers/policies (inference/v2/model_implementations/{llama_v2,mistral,mixtral,opt,falcon,phi}):

  llama / mistral / internlm (attention biases) : RMSNorm, rotary, gated SiLU MLP, GQA           mixtral : + top-2 MoE MLP (experts TP-sharded)
  opt             : LayerNorm, learned positions (+2), ReLU MLP, biases      falcon / phi : parallel attention + MLP block

Forward of a layer: fused (residual-add +) norm -> QKV GEMM -> RoPE + paged-KV append -> paged attention -> output GEMM ->
tensor-parallel all-reduce -> norm -> MLP GEMMs -> all-reduce (reference: llama_v2/model.py:151-213).
Building blocks come from the module registry (`modules/heuristics.py` picks the implementation: 16-bit or fp8 linear, ...).
Weights live in layer containers (`layer_container_base.py`); a container parameter is complete once its checkpoint tensors have
arrived and is then sharded for this tensor-parallel rank (`sharding/`: column slices of qkv/up/gate, row slices of o/down,
vocabulary-parallel unembedding) and re-laid-out by the module that will consume it (`transform_*_param`)."""
import math
import re
from dataclasses import dataclass, field
from typing import Optional

import os

import torch

from .... import comm as dist
from ..inference_utils import ActivationType, DtypeEnum, NormTypeEnum
from ..modules import heuristics
from ..modules.configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig, DSUnembedConfig,
                               PositionalEmbeddingType, RotateHalfConfig)
from . import sharding
from .common_parameters import (AttentionOutputParameter, EmbeddingParameter, GatedMLPParameter, MLP1Parameter, MLP2Parameter,
                                MoEGatingWeightParameter, NormParameter, UnembedParameter, UnfusedMoEGatedMLPParameter, UnfusedMoEMLP2Parameter,
                                UnfusedQKVParameter)
from .layer_container_base import LayerContainer


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


# ---- containers addressed by the canonical parameter names (every family's checkpoint names are mapped onto these sources,
#      either by its own container class under model_implementations/<family>/ or by engine_factory's name mappers)
class _AttnNormContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    qkv_b: Optional[UnfusedQKVParameter]
    o_w: AttentionOutputParameter
    o_b: Optional[AttentionOutputParameter]
    norm1_w: NormParameter
    norm1_b: Optional[NormParameter]
    norm2_w: Optional[NormParameter]
    norm2_b: Optional[NormParameter]

    PARAM_MAPPING = {
        "q.weight": "qkv_w.q_params", "k.weight": "qkv_w.k_params", "v.weight": "qkv_w.v_params",
        "q.bias": "qkv_b.q_params", "k.bias": "qkv_b.k_params", "v.bias": "qkv_b.v_params",
        "o.weight": "o_w.params", "o.bias": "o_b.params",
        "norm1.weight": "norm1_w.params", "norm1.bias": "norm1_b.params",
        "norm2.weight": "norm2_w.params", "norm2.bias": "norm2_b.params",
    }


class GatedLayerContainer(_AttnNormContainer):
    mlp1_w: GatedMLPParameter
    mlp2_w: MLP2Parameter
    PARAM_MAPPING = {"gate.weight": "mlp1_w.gate_params", "up.weight": "mlp1_w.up_params", "down.weight": "mlp2_w.params"}


class PlainLayerContainer(_AttnNormContainer):
    mlp1_w: MLP1Parameter
    mlp1_b: Optional[MLP1Parameter]
    mlp2_w: MLP2Parameter
    mlp2_b: Optional[MLP2Parameter]
    PARAM_MAPPING = {"up.weight": "mlp1_w.params", "up.bias": "mlp1_b.params", "down.weight": "mlp2_w.params", "down.bias": "mlp2_b.params"}


class MoELayerContainer(_AttnNormContainer):
    router_w: MoEGatingWeightParameter
    experts_gu: UnfusedMoEGatedMLPParameter
    experts_dn: UnfusedMoEMLP2Parameter
    PARAM_MAPPING = {"router.weight": "router_w.params", "experts.*.gate.weight": "experts_gu.gating_experts",
                     "experts.*.up.weight": "experts_gu.up_experts", "experts.*.down.weight": "experts_dn.experts"}


class NonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    pos_embed: Optional[EmbeddingParameter]
    final_norm_w: NormParameter
    final_norm_b: Optional[NormParameter]
    lm_head: UnembedParameter
    lm_head_b: Optional[UnembedParameter]
    PARAM_MAPPING = {"embed.weight": "embed.params", "pos_embed.weight": "pos_embed.params", "final_norm.weight": "final_norm_w.params",
                     "final_norm.bias": "final_norm_b.params", "lm_head.weight": "lm_head.params", "lm_head.bias": "lm_head_b.params"}


def layer_container_class(spec: ArchSpec):
    if spec.num_experts:
        return MoELayerContainer
    return GatedLayerContainer if spec.gated_mlp else PlainLayerContainer


_ACT = {("silu", True): ActivationType.SiGLU, ("gelu", True): ActivationType.GEGLU, ("relu", True): ActivationType.ReGLU,
        ("silu", False): ActivationType.SILU, ("gelu", False): ActivationType.GELU, ("relu", False): ActivationType.RELU}


class RaggedTransformer:
    """Holds the TP shard of the weights and runs ragged batches against a DSStateManager."""

    def __init__(self, spec: ArchSpec, state_dict=None, tp_group=None, dtype=torch.bfloat16, device=None, symm=None, quantization_mode="",
                 engine_config=None, layer_container_cls=None, non_transformer_container_cls=None):
        self.spec = spec
        self.quantization_mode = quantization_mode or ""
        self.engine_config = engine_config
        self.tp_group = tp_group
        self.tp = dist.get_world_size(tp_group) if (tp_group is not None and dist.is_initialized()) else 1
        self._graphs, self._graph_warm, self._graph_pool = {}, {}, None  # decode-step CUDA graphs by batch size (forward_graphed)
        self._graphs_on = os.environ.get("DSB200_FASTGEN_GRAPHS", "1") != "0"
        self.rank = dist.get_rank(tp_group) if self.tp > 1 else 0
        self.dtype = dtype
        self.device = device or torch.device("cpu")
        self.symm = symm
        s = spec
        assert s.num_heads % self.tp == 0, "num_heads must be divisible by tp_size"
        self.hq, self.hkv = sharding.get_local_heads(self.rank, self.tp, s.num_heads, s.num_kv_heads)
        self.kv_replicated = s.num_kv_heads < self.tp
        self.vocab_padded = -(-s.vocab_size // self.tp) * self.tp
        self._build_modules()
        self.layers = [(layer_container_cls or layer_container_class(s))(self) for _ in range(s.num_layers)]
        self.non_transformer = (non_transformer_container_cls or NonTransformerContainer)(self)
        if state_dict is not None:
            self.load_canonical(state_dict)

    # ---- facts the parameter classes ask for
    @property
    def head_size(self):
        return self.spec.head_dim

    @property
    def n_heads_q(self):
        return self.spec.num_heads

    @property
    def n_heads_kv(self):
        return self.spec.num_kv_heads

    @property
    def n_experts(self):
        return self.spec.num_experts

    @property
    def tp_rank(self):
        return self.rank

    @property
    def tp_size(self):
        return self.tp

    # convenience views of the non-transformer parameters
    embed = property(lambda self: self.non_transformer.embed)
    pos_embed = property(lambda self: self.non_transformer.get("pos_embed"))
    final_norm_w = property(lambda self: self.non_transformer.final_norm_w)
    final_norm_b = property(lambda self: self.non_transformer.get("final_norm_b"))
    lm_head = property(lambda self: self.non_transformer.lm_head)
    lm_head_b = property(lambda self: self.non_transformer.get("lm_head_b"))

    # ---- building blocks (modules/heuristics.py chooses the implementation of each)
    def _build_modules(self):
        s, ec, dt = self.spec, self.engine_config, DtypeEnum(self.dtype)
        sm = getattr(ec, "state_manager", None)
        max_tokens = getattr(sm, "max_ragged_batch_size", 768)
        max_seqs = getattr(sm, "max_ragged_sequence_count", 512)
        if ec is None and self.quantization_mode:
            from ..config_v2 import QuantizationConfig, RaggedInferenceEngineConfig
            ec = RaggedInferenceEngineConfig(quantization=QuantizationConfig(quantization_mode=self.quantization_mode))
        H, d = s.hidden_size, s.head_dim
        f_lo, f_hi = sharding.get_shard_endpoints(s.intermediate_size, self.rank, self.tp, self._mlp_granularity())
        norm_type = NormTypeEnum.RMSNorm if s.norm == "rms" else NormTypeEnum.LayerNorm
        act = _ACT[(s.activation, bool(s.gated_mlp or s.num_experts))]
        common = dict(input_dtype=dt, output_dtype=dt)
        self.embed_mod = heuristics.instantiate_embed(DSEmbeddingsConfig(max_tokens=max_tokens, residual_dtype=dt, embedding_dim=H,
                                                                         positional_embedding=s.learned_pos_offset is not None,
                                                                         positional_offset=s.learned_pos_offset or 0), ec)
        self.norm_mod = heuristics.instantiate_pre_norm(DSNormConfig(max_tokens=max_tokens, type=norm_type, channels=H, residual_dtype=dt,
                                                                      eps=s.norm_eps, **common), ec)
        rotary = dict(positional_embedding_type=PositionalEmbeddingType.rotate_half,
                      positional_embedding_config=RotateHalfConfig(theta_base=s.rope_theta, rotate_dim=s.rotary_dim)) if s.rotary_dim else {}
        self.attn_mod = heuristics.instantiate_attention(DSSelfAttentionConfig(max_tokens=max_tokens, n_heads_q=self.hq, n_heads_kv=self.hkv,
                                                                               head_size=d, max_sequences=max_seqs, **common, **rotary), ec)
        self.qkv_lin = heuristics.instantiate_linear(DSLinearConfig(max_tokens=max_tokens, in_channels=H, out_channels=(self.hq + 2 * self.hkv) * d,
                                                                    **common), ec)
        self.o_lin = heuristics.instantiate_linear(DSLinearConfig(max_tokens=max_tokens, in_channels=self.hq * d, out_channels=H, **common), ec)
        self.mlp1_lin = self.mlp2_lin = self.moe_mod = None
        if s.num_experts:
            self.moe_mod = heuristics.instantiate_moe(DSMoEConfig(max_tokens=max_tokens, model_dim=H, intermediate_features=f_hi - f_lo,
                                                                  n_experts=s.num_experts, top_k=s.top_k, activation=act, normalize_scores=True,
                                                                  **common), ec)
        else:
            self.mlp1_lin = heuristics.instantiate_linear(DSLinearConfig(max_tokens=max_tokens, in_channels=H, out_channels=f_hi - f_lo,
                                                                         activation=act, **common), ec)
            self.mlp2_lin = heuristics.instantiate_linear(DSLinearConfig(max_tokens=max_tokens, in_channels=f_hi - f_lo, out_channels=H, **common), ec)
        self.unembed_mod = heuristics.instantiate_unembed(DSUnembedConfig(max_tokens=max_tokens, dtype=dt, norm_type=norm_type, model_dim=H,
                                                                          max_sequences=max_seqs, vocab_size=self.vocab_padded // self.tp,
                                                                          eps=s.norm_eps), ec)

    def _mlp_granularity(self):
        g = sharding.DEFAULT_SHARD_GRANULARITY
        return g if self.spec.intermediate_size % g == 0 else 1

    # ---- parameter transforms: shard for this rank, then let the consuming module lay the tensor out
    def _dev(self, t):
        return t.to(self.device)

    def transform_embedding_param(self, param):
        return self._dev(self.embed_mod.transform_param(param))  # whole table on every rank: one gather launch, no collective

    def transform_qkv_param(self, param):
        s = self.spec
        local = sharding.shard_qkv_param(param, self.rank, self.tp, s.head_dim, s.num_heads, s.num_kv_heads)
        return self._dev(self.qkv_lin.transform_param(local))

    def transform_attn_out_param(self, param):
        s = self.spec
        if param.ndim == 1:  # the bias is added after the all-reduce, on every rank
            return self._dev(param.to(self.dtype).contiguous())
        local = sharding.shard_attn_out_param(param, self.rank, self.tp, s.head_dim, s.num_heads, s.num_kv_heads)
        return self._dev(self.o_lin.transform_param(local))

    def transform_mlp_1_param(self, param, gated=False):
        local = sharding.shard_param(param, sharding.ShardingType.OUTER_DIMENSION, self.rank, self.tp,
                                     num_concatenated_matrices=2 if gated else 1, granularity=self._mlp_granularity())
        return self._dev(self.mlp1_lin.transform_param(local))

    def transform_mlp_2_param(self, param):
        if param.ndim == 1:
            return self._dev(param.to(self.dtype).contiguous())
        local = sharding.shard_param(param, sharding.ShardingType.INNER_DIMENSION, self.rank, self.tp, granularity=self._mlp_granularity())
        return self._dev(self.mlp2_lin.transform_param(local))

    def transform_norm_param(self, param):
        return self._dev(self.norm_mod.transform_param(param))

    def transform_inv_freq_param(self, param):
        return self._dev(param.to(torch.float32).contiguous())

    def transform_unembed_param(self, param):
        pad = self.vocab_padded - param.shape[0]
        if pad:  # equal vocabulary ranges on every rank (the logits all-gather wants one size)
            param = torch.cat([param, param.new_zeros(pad, *param.shape[1:])])
        return self._dev(sharding.shard_unembed_param(param, self.rank, self.tp).to(self.dtype).contiguous())

    def transform_moe_gate_param(self, param):
        return self._dev(self.moe_mod.transform_gate_param(param))

    def transform_moe_mlp_1_param(self, param, gated=True):
        local = sharding.shard_param(param, sharding.ShardingType.OUTER_DIMENSION, self.rank, self.tp,
                                     num_concatenated_matrices=2 if gated else 1, granularity=self._mlp_granularity(), bias_dims=2)
        return self._dev(self.moe_mod.transform_moe_mlp_1_param(local))

    def transform_moe_mlp_2_param(self, param):
        local = sharding.shard_param(param, sharding.ShardingType.INNER_DIMENSION, self.rank, self.tp, granularity=self._mlp_granularity())
        return self._dev(self.moe_mod.transform_moe_mlp_2_param(local))

    # ---- loading
    _LAYER_RX = re.compile(r"^layers\.(\d+)\.(.*)$")

    def set_dependency(self, name: str, tensor: torch.Tensor) -> bool:
        """Route one canonically named checkpoint tensor to the container that owns it."""
        m = self._LAYER_RX.match(name)
        if m:
            return self.layers[int(m.group(1))].set_dependency(m.group(2), tensor)
        return self.non_transformer.set_dependency(name, tensor)

    def load_canonical(self, sd):
        for name, tensor in sd.items():
            if not self.set_dependency(name, tensor):
                raise KeyError(f"unexpected parameter '{name}' for family {self.spec.family}")
        if self.spec.tie_embeddings or "lm_head.weight" not in sd:
            self.non_transformer.set_dependency("lm_head.weight", sd["embed.weight"])
        self.check_populated()

    def check_populated(self):
        for i, c in enumerate(list(self.layers) + [self.non_transformer]):
            if not c.is_populated:
                where = f"layer {i}" if i < len(self.layers) else "non-transformer parameters"
                raise RuntimeError(f"{where}: missing checkpoint tensors for {c.missing()}")

    def named_containers(self):
        for i, c in enumerate(self.layers):
            yield f"layers.{i}", c
        yield "non_transformer", self.non_transformer

    # ---- forward
    def _tp_ops(self):
        """Peer-memory TP collectives (fused GEMM + all-reduce, switch all-reduce, logits push), built on first use; None when
        the group is not a single-node CUDA bf16 group (then NCCL is used)."""
        ops = getattr(self, "_tp_coll", False)
        if ops is False:
            ops = None
            import os
            if (self.tp > 1 and self.dtype == torch.bfloat16 and torch.device(self.device).type == "cuda" and
                    os.environ.get("DSB200_DISABLE_P2P", "0") != "1" and self.tp <= torch.cuda.device_count()):
                try:
                    from ..tp_collectives import RaggedTPCollectives
                    sm = getattr(self.engine_config, "state_manager", None)
                    ops = RaggedTPCollectives(self.tp_group, self.spec.hidden_size, getattr(sm, "max_ragged_batch_size", 768),
                                              getattr(sm, "max_ragged_sequence_count", 512), self.vocab_padded // self.tp)
                except Exception as e:  # no peer access: stay on NCCL
                    from ....utils.logging import logger
                    logger.warning(f"FastGen peer-memory TP collectives unavailable ({e}); using NCCL")
                    ops = None
            self._tp_coll = ops
        return ops

    def _all_reduce(self, x):
        if self.tp > 1:
            ops = self._tp_ops()
            if ops is not None:
                return ops.all_reduce(x if x.is_contiguous() else x.contiguous())
            dist.all_reduce(x, group=self.tp_group)
        return x

    def _row_parallel(self, lin, x, w):
        """all-reduce(lin(x, w)): one fused tcgen05 GEMM + peer all-reduce kernel when the weight is plain bf16."""
        ops = self._tp_ops() if self.tp > 1 else None
        if ops is not None and ops.usable_for(x, w):
            return ops.row_parallel(x, w)
        return self._all_reduce(lin(x, w))

    def _mlp(self, L, h):
        if self.moe_mod is not None:
            return self.moe_mod(h, L["router_w"], L["experts_gu"], L["experts_dn"])
        return self.mlp2_lin(self.mlp1_lin(h, L["mlp1_w"], L.get("mlp1_b")), L["mlp2_w"])

    def forward_graphed(self, batch, state_manager):
        """`forward`, with pure decode steps (one new token per sequence) replayed from a CUDA graph per batch size.

        A decode step of a deep model is ~12 small launches per layer; issued one by one from Python they cost more host time than
        the GPU needs to run them.  Everything a step reads sits at fixed device addresses (the ragged batch's metadata buffers,
        the KV cache, the weights, the symmetric buffers; the epochs of the peer-memory collectives live in device memory), so
        the launch sequence of a step with n sequences is captured once -- after two eager steps of that size -- and replayed.
        Prefill / mixed batches run eagerly.  DSB200_FASTGEN_GRAPHS=0 turns this off."""
        n = batch.current_sequences
        if not (self._graphs_on and n == batch.current_tokens and n > 0 and batch.input_ids().is_cuda):
            return self.forward(batch, state_manager)
        entry = self._graphs.get(n)
        if entry is None:
            seen = self._graph_warm.get(n, 0)
            if seen < 2:  # lazily built state (tensor maps, split heuristics, module buffers) settles in eager mode first
                self._graph_warm[n] = seen + 1
                return self.forward(batch, state_manager)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g, pool=self._graph_pool):
                    out = self.forward(batch, state_manager)
            except Exception as e:  # something on this model's path is not capturable: stay eager from now on
                from ....utils.logging import logger
                logger.warning(f"FastGen decode step could not be captured into a CUDA graph ({type(e).__name__}: {e}); running eagerly")
                self._graphs_on = False
                torch.cuda.synchronize()
                return self.forward(batch, state_manager)
            if self._graph_pool is None:
                self._graph_pool = g.pool()
            entry = self._graphs[n] = (g, out)
        g, out = entry
        g.replay()
        return out.clone()  # the graph's output buffer is rewritten by the next replay

    def forward(self, batch, state_manager):
        """batch: RaggedBatchWrapper (finalized).  Returns logits [n_seqs, vocab] for the last token of every sequence."""
        s, nt = self.spec, self.non_transformer
        x = self.embed_mod(batch, nt.embed, nt.get("pos_embed"))
        residual = None
        for li, L in enumerate(self.layers):
            # residual += x (the previous layer's MLP output), hn = norm(residual)
            residual, hn = self.norm_mod(x if residual is None else residual, None if residual is None else x, L["norm1_w"], L.get("norm1_b"))
            qkv = self.qkv_lin(hn, L["qkv_w"], L.get("qkv_b"))
            attn = self.attn_mod(qkv, state_manager.get_cache(li), batch)
            if not s.parallel_block:
                attn_out = self._row_parallel(self.o_lin, attn, L["o_w"])
            else:
                attn_out = self.o_lin(attn, L["o_w"])
            if s.parallel_block:
                # gpt-neox gives the MLP its own LayerNorm of the block input (= the residual stream after the fused add)
                mlp_in = self.norm_mod(residual, None, L["norm2_w"], L.get("norm2_b"))[1] if s.parallel_norm2 else hn
                x = self._all_reduce(attn_out + self._mlp(L, mlp_in))
                if "o_b" in L:
                    x = x + L["o_b"]
                if "mlp2_b" in L:
                    x = x + L["mlp2_b"]
                continue
            if "o_b" in L:
                attn_out = attn_out + L["o_b"]
            residual, hn = self.norm_mod(residual, attn_out, L["norm2_w"], L.get("norm2_b"))
            if self.moe_mod is None:
                x = self._row_parallel(self.mlp2_lin, self.mlp1_lin(hn, L["mlp1_w"], L.get("mlp1_b")), L["mlp2_w"])
            else:
                x = self._all_reduce(self._mlp(L, hn))
            if "mlp2_b" in L:
                x = x + L["mlp2_b"]
        logits = self.unembed_mod(x, nt.lm_head, batch, bias=nt.get("lm_head_b"), gamma=nt.final_norm_w, beta=nt.get("final_norm_b"),
                                  residual=residual)
        if self.tp > 1:
            ops = self._tp_ops()
            if ops is not None:
                logits = ops.gather_logits(logits)
            else:
                full = torch.empty(self.tp, *logits.shape, dtype=logits.dtype, device=logits.device)
                dist.all_gather_into_tensor(full, logits.contiguous(), group=self.tp_group)
                logits = full.permute(1, 0, 2).reshape(logits.shape[0], -1)
        return logits[:, :s.vocab_size]
