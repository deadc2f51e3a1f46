"""FastGen building blocks below the engine: module registry + heuristics, sharding arithmetic, declarative parameters and
containers, policies, flattened-parameter serialisation (reference test strategy: tests/unit/inference/v2/{modules,
model_implementations/{parameters,sharding}}/*)."""
import pytest
import torch

from test_inference_v2 import CFG, _hf_llama_state

from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict, build_hf_engine
from deepspeedsyclsupport_b200.inference.v2.inference_parameter import InferenceParameter
from deepspeedsyclsupport_b200.inference.v2.inference_utils import ActivationType, DtypeEnum, NormTypeEnum, ceil_div, elem_size, is_gated
from deepspeedsyclsupport_b200.inference.v2.model_implementations import sharding
from deepspeedsyclsupport_b200.inference.v2.model_implementations.common_parameters import (GQAMegatronQKVParameter, MegatronQKVParameter,
                                                                                           UnfusedMoEGatedMLPParameter, UnfusedQKVParameter)
from deepspeedsyclsupport_b200.inference.v2.model_implementations.layer_container_base import LayerContainer
from deepspeedsyclsupport_b200.inference.v2.model_implementations.parameter_base import ParameterBase, ParamList
from deepspeedsyclsupport_b200.inference.v2.modules import heuristics
from deepspeedsyclsupport_b200.inference.v2.modules.configs import DSLinearConfig, DSNormConfig
from deepspeedsyclsupport_b200.inference.v2.modules.interfaces import DSLinearBase, DSLinearRegistry, DSPreNormRegistry
from deepspeedsyclsupport_b200.inference.v2.modules.module_registry import ConfigBundle


def _ecfg(mode=""):
    return RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64, "max_ragged_sequence_count": 8, "max_context": 128,
                                                      "memory_config": {"mode": "allocate", "size": 40}},
                                       quantization={"quantization_mode": mode})


# ------------------------------------------------------------------ utils
def test_enums_and_helpers():
    assert DtypeEnum("bf16") is DtypeEnum.bf16 is DtypeEnum(torch.bfloat16) is DtypeEnum("torch.bfloat16")
    assert DtypeEnum("half").value is torch.float16
    assert is_gated(ActivationType.SiGLU) and is_gated(3) and not is_gated(ActivationType.GELU)
    assert elem_size(torch.bfloat16) == 2 and elem_size(torch.float8_e4m3fn) == 1 and ceil_div(7, 3) == 3
    with pytest.raises(ValueError):
        elem_size("nope")


def test_inference_parameter_keeps_aux_tensors():
    p = InferenceParameter.initialize(torch.randn(4, 8), scales=torch.arange(4.0))
    q = p.to(torch.bfloat16)
    assert isinstance(q, InferenceParameter) and q.dtype == torch.bfloat16 and q.scales.dtype == torch.float32
    assert type(p + 1) is torch.Tensor  # operations leave the subclass: no dispatch overhead on the forward path
    with pytest.raises(ValueError):
        InferenceParameter.initialize(torch.zeros(2), shape=torch.zeros(1))  # would shadow a tensor attribute
    with pytest.raises(ValueError):
        InferenceParameter.initialize(torch.zeros(2), scale=1.0)


# ------------------------------------------------------------------ registry / heuristics
def test_registry_rejects_foreign_classes_and_unknown_names():
    with pytest.raises(TypeError):
        DSLinearRegistry.register_module(torch.nn.Linear)
    cfg = DSLinearConfig(max_tokens=8, in_channels=16, out_channels=16, input_dtype="fp32", output_dtype="fp32")
    with pytest.raises(KeyError):
        DSLinearRegistry.instantiate_config(ConfigBundle(name="no_such_linear", config=cfg))
    mixed = DSLinearConfig(max_tokens=8, in_channels=16, out_channels=16, input_dtype="fp16", output_dtype="bf16")
    with pytest.raises(ValueError):
        DSLinearRegistry.instantiate_config(ConfigBundle(name="blas_fp_linear", config=mixed))
    assert {"blas_fp_linear", "fp8_linear"} <= set(DSLinearRegistry.registry) and {"cuda_pre_ln", "cuda_pre_rms"} <= set(DSPreNormRegistry.registry)

    @DSLinearRegistry.register_module
    class Doubling(DSLinearBase):  # a user-provided implementation joins the registry and is instantiable by name

        @staticmethod
        def name():
            return "test_doubling_linear"

        @staticmethod
        def supports_config(config):
            return True

        def transform_param(self, param):
            return InferenceParameter.initialize(param)

        def forward(self, hidden_states, w, b=None):
            return 2 * torch.nn.functional.linear(hidden_states, w, b)

    mod = DSLinearRegistry.instantiate_config(ConfigBundle(name="test_doubling_linear", config=cfg))
    x, w = torch.randn(3, 16), torch.randn(16, 16)
    assert torch.allclose(mod(x, w), 2 * x @ w.t(), atol=1e-5)
    del DSLinearRegistry.registry["test_doubling_linear"]


def test_heuristics_follow_engine_config():
    cfg = DSLinearConfig(max_tokens=8, in_channels=32, out_channels=16, input_dtype="fp32", output_dtype="fp32")
    assert type(heuristics.instantiate_linear(cfg, _ecfg())).__name__ == "BlasFPLinear"
    assert type(heuristics.instantiate_linear(cfg, _ecfg("wf8af8"))).__name__ == "FP8Linear"
    odd = DSLinearConfig(max_tokens=8, in_channels=24, out_channels=16, input_dtype="fp32", output_dtype="fp32")
    assert type(heuristics.instantiate_linear(odd, _ecfg("wf8af8"))).__name__ == "BlasFPLinear"  # K not a multiple of 16: stays 16-bit
    rms = DSNormConfig(max_tokens=8, type=NormTypeEnum.RMSNorm, channels=32, input_dtype="fp32", output_dtype="fp32", residual_dtype="fp32")
    ln = DSNormConfig(max_tokens=8, type="layer_norm", channels=32, input_dtype="fp32", output_dtype="fp32", residual_dtype="fp32")
    assert heuristics.instantiate_pre_norm(rms).name() == "cuda_pre_rms" and heuristics.instantiate_pre_norm(ln).name() == "cuda_pre_ln"
    x, h, g, b = torch.randn(5, 32), torch.randn(5, 32), torch.rand(32) + 0.5, torch.randn(32)
    res, out = heuristics.instantiate_pre_norm(ln)(x, h, g, b)
    assert torch.allclose(res, x + h) and torch.allclose(out, torch.nn.functional.layer_norm(x + h, (32, ), g, b), atol=1e-5)
    res, out = heuristics.instantiate_post_norm(ln)(x, h, g, b)
    assert torch.allclose(res, torch.nn.functional.layer_norm(x + h, (32, ), g, b), atol=1e-5) and res is out


@pytest.mark.parametrize("act,gated", [(ActivationType.GELU, False), (ActivationType.RELU, False), (ActivationType.GEGLU, True),
                                       (ActivationType.IDENTITY, False)])
def test_linear_module_activation_epilogues(act, gated):
    import torch.nn.functional as F
    cfg = DSLinearConfig(max_tokens=8, in_channels=32, out_channels=24, activation=act, input_dtype="fp32", output_dtype="fp32")
    lin = heuristics.instantiate_linear(cfg)
    w, b, x = lin.transform_param(torch.randn(48 if gated else 24, 32)), torch.randn(48 if gated else 24), torch.randn(6, 32)
    y = F.linear(x, w, b)
    fn = {ActivationType.GELU: lambda t: F.gelu(t, approximate="tanh"), ActivationType.RELU: F.relu, ActivationType.GEGLU: lambda t: F.gelu(
        t, approximate="tanh"), ActivationType.IDENTITY: lambda t: t}[act]
    want = fn(y[:, :24]) * y[:, 24:] if gated else fn(y)
    assert torch.allclose(lin(x, w, b), want, atol=1e-5) and lin.output is not None


# ------------------------------------------------------------------ sharding
def test_shard_endpoints_spread_the_remainder():
    assert [sharding.get_shard_endpoints(96, r, 4, 32) for r in range(4)] == [(0, 32), (32, 64), (64, 96), (96, 96)]
    assert [sharding.get_shard_endpoints(160, r, 2, 32) for r in range(2)] == [(0, 96), (96, 160)]
    sizes = [sharding.sharded_unembed_dim(50257, r, 8) for r in range(8)]
    assert sum(sizes) == 50257 and max(sizes) - min(sizes) == 1
    with pytest.raises(AssertionError):
        sharding.get_shard_endpoints(100, 0, 2, 32)


@pytest.mark.parametrize("hq,hkv", [(8, 8), (8, 2), (8, 1), (12, 4)])
@pytest.mark.parametrize("world", [1, 2, 4])
def test_qkv_and_attn_out_shards_reassemble(hq, hkv, world):
    d, hidden = 4, 24
    if hq % world or (hkv >= world and hkv % world) or (hkv < world and world % hkv):
        pytest.skip("not a valid split")
    w = torch.arange((hq + 2 * hkv) * d * hidden, dtype=torch.float32).view(-1, hidden)
    o = torch.randn(hidden, hq * d)
    q_rows, o_cols, group = [], [], hq // hkv
    for r in range(world):
        s = sharding.shard_qkv_param(w, r, world, d, hq, hkv)
        lq, lkv = sharding.get_local_heads(r, world, hq, hkv)
        assert s.shape[0] == (lq + 2 * lkv) * d == sharding.qkv_out_features(hq * d, r, world, d, hq, hkv)
        q_rows.append(s[:lq * d])
        # the K head(s) on this rank are the ones its query heads belong to
        first_q_head = sum(sharding.get_local_heads(i, world, hq, hkv)[0] for i in range(r))
        kv_head = first_q_head // group
        assert torch.equal(s[lq * d:lq * d + d], w[(hq + kv_head) * d:(hq + kv_head + 1) * d])
        oc = sharding.shard_attn_out_param(o, r, world, d, hq, hkv)
        assert oc.shape[1] == sharding.attn_out_in_features(hidden, r, world, d, hq, hkv) == lq * d
        o_cols.append(oc)
    assert torch.equal(torch.cat(q_rows), w[:hq * d]) and torch.equal(torch.cat(o_cols, 1), o)
    # a bias follows the weight's rows; the row-parallel bias lives on rank 0 only
    bias = torch.arange((hq + 2 * hkv) * d, dtype=torch.float32)
    assert torch.equal(sharding.shard_qkv_param(bias, 0, world, d, hq, hkv), sharding.shard_qkv_param(w, 0, world, d, hq, hkv)[:, 0] / hidden)
    assert sharding.shard_attn_out_param(torch.zeros(hidden), 0, world, d, hq, hkv) is not None
    if world > 1:
        assert sharding.shard_attn_out_param(torch.zeros(hidden), 1, world, d, hq, hkv) is None


def test_gated_mlp_and_moe_shards_cut_both_halves_alike():
    inter, hidden, world = 128, 16, 4
    gate_up = torch.randn(2 * inter, hidden)
    down = torch.randn(hidden, inter)
    x = torch.randn(3, hidden)
    full = (torch.nn.functional.silu(x @ gate_up[:inter].t()) * (x @ gate_up[inter:].t())) @ down.t()
    acc = 0
    for r in range(world):
        gu = sharding.shard_mlp_1_param(gate_up, r, world, gated=True)
        dn = sharding.shard_mlp_2_param(down, r, world)
        assert gu.shape[0] == 2 * sharding.sharded_intermediate_dim(inter, world, r)
        h = x @ gu.t()
        acc = acc + (torch.nn.functional.silu(h[:, :gu.shape[0] // 2]) * h[:, gu.shape[0] // 2:]) @ dn.t()
    assert torch.allclose(acc, full, atol=1e-4)
    experts = torch.randn(3, 2 * inter, hidden)
    assert sharding.shard_mlp_1_param(experts, 1, world, gated=True, is_moe=True).shape == (3, 2 * inter // world, hidden)
    expert_bias = torch.randn(3, inter)
    assert sharding.shard_mlp_1_param(expert_bias, 1, world, is_moe=True).shape == (3, inter // world)
    assert sharding.shard_mlp_2_param(torch.randn(3, hidden), 1, world, is_moe=True) is None


# ------------------------------------------------------------------ parameters / containers
class _FakeModel:
    n_experts, head_size, n_heads_q, n_heads_kv = 3, 2, 4, 2

    def transform_qkv_param(self, p):
        return p.clone()

    def transform_moe_mlp_1_param(self, p, gated=True):
        return p.clone()


def test_parameter_finalizes_when_last_source_arrives():
    done = []
    p = UnfusedQKVParameter(_FakeModel(), on_complete=done.append)
    p.q_params, p.k_params = torch.ones(4, 2), 2 * torch.ones(2, 2)
    assert not p.is_finalized and p.missing() == ["v_params"] and p.touched
    with pytest.raises(ValueError):
        p.q_params = torch.zeros(4, 2)
    with pytest.raises(RuntimeError):
        _ = p.value
    p.v_params = 3 * torch.ones(2, 2)
    assert p.is_finalized and done == [p] and p.value.shape == (8, 2) and p.value[:, 0].tolist() == [1, 1, 1, 1, 2, 2, 3, 3]
    m = UnfusedMoEGatedMLPParameter(_FakeModel())
    for e in (2, 0, 1):  # experts may arrive in any order
        m.gating_experts[e] = torch.full((2, 2), float(e))
        m.up_experts[e] = torch.full((2, 2), 10.0 + e)
    assert m.value.shape == (3, 4, 2) and m.value[1, :, 0].tolist() == [1, 1, 11, 11]
    with pytest.raises(AttributeError):
        m.gating_experts = []


def test_interleaved_qkv_layouts_become_q_k_v():
    model = _FakeModel()
    hq, hkv, d = 4, 2, 2
    q, k, v = torch.arange(8.).view(8, 1), 100 + torch.arange(4.).view(4, 1), 200 + torch.arange(4.).view(4, 1)
    mha = torch.stack([torch.stack([q[h * d:(h + 1) * d], 100 + q[h * d:(h + 1) * d], 200 + q[h * d:(h + 1) * d]]) for h in range(hq)]).reshape(-1, 1)
    model.n_heads_kv = 4
    p = MegatronQKVParameter(model)
    p.params = mha
    assert torch.equal(p.value, torch.cat([q, 100 + q, 200 + q]))
    model.n_heads_kv = 2
    grouped = torch.cat([torch.cat([q[g * 4:(g + 1) * 4], k[g * 2:(g + 1) * 2], v[g * 2:(g + 1) * 2]]) for g in range(hkv)])
    p = GQAMegatronQKVParameter(model)
    p.params = grouped
    assert torch.equal(p.value, torch.cat([q, k, v]))


def test_layer_container_routes_names_wildcards_and_transforms():
    from typing import Optional

    class Doubled(ParameterBase):
        params: torch.Tensor

        def finalize(self):
            return 2 * self.params

    class Stack(ParameterBase):
        experts: ParamList("n_experts")

        def finalize(self):
            return torch.stack(list(self.experts))

    class C(LayerContainer):
        a: Doubled
        b: Optional[Doubled]
        e: Stack
        PARAM_MAPPING = {"x.weight": ["a.params", "b.params"], "experts.*.w": ("e.experts", "neg")}

        @staticmethod
        def neg(t):
            return -t

    class D(C):  # subclasses extend the mapping
        c: Optional[Doubled]
        PARAM_MAPPING = {"y.weight": "c.params"}

    c = D(_FakeModel())
    assert not c.is_populated and "a.params" in c.missing()
    assert not c.set_dependency("unknown.weight", torch.zeros(1))
    assert c.set_dependency("x.weight", torch.ones(2))
    for i in range(3):
        assert c.set_dependency(f"experts.{i}.w", torch.full((2, ), float(i)))
    assert c.is_populated and c.a.tolist() == [2, 2] and c["b"].tolist() == [2, 2] and c.e[2].tolist() == [-2, -2]
    assert "c" not in c and c.get("c") is None and c.c is None and [n for n, _ in c.items()] == ["a", "b", "e"]
    with pytest.raises(KeyError):
        c["c"]
    with pytest.raises(ValueError):

        class Bad(LayerContainer):
            a: Doubled
            PARAM_MAPPING = {"x": "a.nope"}


# ------------------------------------------------------------------ policies / serialisation
def test_policy_reports_unknown_and_missing_tensors():
    from deepspeedsyclsupport_b200.inference.v2.checkpoint import InMemoryModelEngine
    from deepspeedsyclsupport_b200.inference.v2.engine_factory import build_model_from_checkpoint
    sd = _hf_llama_state(CFG)
    with pytest.raises(ValueError, match="Cannot find container"):
        build_model_from_checkpoint("llama", CFG, InMemoryModelEngine(dict(sd, **{"model.layers.0.mystery.weight": torch.zeros(1)})), _ecfg(),
                                    dtype=torch.float32)
    short = {k: v for k, v in sd.items() if k != "model.layers.1.mlp.up_proj.weight"}
    with pytest.raises(ValueError, match="mlp1_w.up_params"):
        build_model_from_checkpoint("llama", CFG, InMemoryModelEngine(short), _ecfg(), dtype=torch.float32)
    with pytest.raises(ValueError, match="Unsupported model type"):
        build_engine_from_state_dict("not_a_model", CFG, sd, _ecfg())
    tied = {k: v for k, v in sd.items() if k != "lm_head.weight"}  # tied embeddings: the table also serves as unembedding
    m = build_model_from_checkpoint("llama", CFG, InMemoryModelEngine(tied), _ecfg(), dtype=torch.float32)
    assert torch.equal(m.lm_head, sd["model.embed_tokens.weight"])
    assert m.num_layers == 2 and m.model_dim == 64 and m.n_heads_q_local == 4 and type(m).__name__ == "Llama2InferenceModel"


@pytest.mark.parametrize("mode", ["", "wf8af8"])
def test_serialize_writes_one_flat_buffer_and_restores_identically(tmp_path, mode):
    import json
    import os
    sd = _hf_llama_state(CFG)
    eng = build_engine_from_state_dict("llama", CFG, sd, _ecfg(mode), dtype=torch.float32)
    ids = torch.randint(0, 97, (19, ), generator=torch.Generator().manual_seed(3))
    before = eng.put([1], [ids]).clone()
    eng.serialize(str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == ["ds_model_config.pkl", "metadata_rank_0_of_1.json", "params_rank_0_of_1.pt"]
    md = json.load(open(tmp_path / "metadata_rank_0_of_1.json"))
    qkv = md["layers"]["layers.0"]["params"]["qkv_w"]
    assert qkv["core_param"]["offset"] % 256 == 0 and (("scales" in qkv["aux_params"]) == (mode == "wf8af8"))
    # after flattening the live parameters alias the buffer; the engine keeps producing the same logits
    eng.flush(1)
    assert torch.equal(eng.put([1], [ids]), before)
    eng2 = build_hf_engine(str(tmp_path), _ecfg(mode))  # detects the serialised form
    assert eng2._model.layers[0]["qkv_w"].dtype == (torch.float8_e4m3fn if mode else torch.float32)
    assert torch.equal(eng2.put([7], [ids]), before)
    base = eng2._model.flattened_params
    p = eng2._model.layers[1]["mlp2_w"]
    assert base.data_ptr() <= p.data_ptr() < base.data_ptr() + base.numel()


# ------------------------------------------------------------------ kernel objects
def test_kernel_objects_validate_at_construction_and_write_into_outputs():
    import torch.nn.functional as F
    from deepspeedsyclsupport_b200.inference.v2.kernels.core_ops import (BlasLibLinear, CUDABiasActivation, CUDAFPPostLN, CUDAFPPreLN, CUDAGatedActivation,
                                                                      CUDARMSNorm, CUDARMSPreNorm)
    from deepspeedsyclsupport_b200.inference.v2.kernels.cutlass_ops import MixedGEMM, MoEGEMM
    from deepspeedsyclsupport_b200.inference.v2.kernels.ds_kernel import DSKernelBase
    x, h, g, b = torch.randn(4, 32), torch.randn(4, 32), torch.rand(32) + 0.5, torch.randn(32)
    for bad in (lambda: CUDARMSNorm(30, "bf16"), lambda: CUDARMSNorm(32, torch.int8), lambda: CUDAGatedActivation(32, "fp32", ActivationType.GELU),
                lambda: CUDABiasActivation(32, "fp32", ActivationType.SiGLU), lambda: MixedGEMM("bf16", ActivationType.IDENTITY, 3),
                lambda: MoEGEMM("fp32", ActivationType.SILU)):
        with pytest.raises(ValueError):
            bad()
    assert issubclass(CUDARMSNorm, DSKernelBase)
    z = torch.empty_like(x)
    rms = lambda t: t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + 1e-5) * g
    assert CUDARMSNorm(32, "fp32")(z, x, g) is z and torch.allclose(z, rms(x), atol=1e-5)
    zr, zh = torch.empty_like(x), torch.empty_like(x)
    CUDARMSPreNorm(32, "fp32")(zr, zh, x, h, g)
    assert torch.allclose(zr, x + h) and torch.allclose(zh, rms(x + h), atol=1e-5)
    CUDAFPPreLN(32, "fp32")(zr, zh, x, h, g, b)
    assert torch.allclose(zh, F.layer_norm(x + h, (32, ), g, b), atol=1e-5)
    CUDAFPPostLN(32, "fp32")(z, x, h, g, b)
    assert torch.allclose(z, F.layer_norm(x + h, (32, ), g, b), atol=1e-5)
    act = x.clone()
    CUDABiasActivation(32, "fp32", ActivationType.RELU)(act, b)
    assert torch.allclose(act, F.relu(x + b))
    out = torch.empty(4, 16)
    CUDAGatedActivation(16, "fp32", ActivationType.SiGLU)(out, x, b)
    assert torch.allclose(out, F.silu((x + b)[:, :16]) * (x + b)[:, 16:], atol=1e-6)
    w = torch.randn(24, 32)
    o = torch.empty(4, 24)
    BlasLibLinear("fp32")(o, x, w)
    assert torch.allclose(o, x @ w.t(), atol=1e-5)
    # weight-only quantised GEMMs: 8-bit and packed 4-bit, one scale per group of 8 input features
    w8, sc = torch.randint(-127, 127, (24, 32), dtype=torch.int8), torch.rand(24, 4) * 0.01
    MixedGEMM("bf16", ActivationType.IDENTITY, 8)(o, x, w8, sc)
    assert torch.allclose(o, x @ (w8.float().view(24, 4, 8) * sc[..., None]).view(24, 32).t(), atol=1e-4)
    q4 = torch.randint(0, 16, (24, 32))
    packed = (q4[:, 0::2] | (q4[:, 1::2] << 4)).to(torch.uint8)
    MixedGEMM("bf16", ActivationType.IDENTITY, 4)(o, x, packed, sc)
    assert torch.allclose(o, x @ ((q4 - 8).float().view(24, 4, 8) * sc[..., None]).view(24, 32).t(), atol=1e-4)
    xs, W, cum, oo = torch.randn(6, 32), torch.randn(2, 8, 32), torch.tensor([4, 6]), torch.empty(6, 8)
    MoEGEMM("bf16", ActivationType.SILU)(oo, xs, W, cum, torch.ones(2, 8))
    assert torch.allclose(oo, F.silu(torch.cat([xs[:4] @ W[0].t(), xs[4:] @ W[1].t()]) + 1), atol=1e-5)


def test_ragged_kernel_objects_against_a_live_batch():
    from deepspeedsyclsupport_b200.inference.v2.kernels.ragged_ops import (AtomBuilder, BlockedFlashAttn, BlockedRotaryEmbeddings, LinearBlockedKVCopy,
                                                                        MoEGather, MoEScatter, RaggedEmbeddingKernel, RaggedLogitsGather,
                                                                        RaggedTopKGating)
    with pytest.raises(ValueError):
        BlockedFlashAttn(72, "bf16")
    with pytest.raises(ValueError):
        RaggedEmbeddingKernel(torch.float32, torch.float32, 64)
    sd = _hf_llama_state(CFG)
    eng = build_engine_from_state_dict("llama", CFG, sd, _ecfg(), dtype=torch.float32)
    a, b = torch.arange(5), torch.arange(10, 13)
    eng.put([1, 2], [a, b])
    batch, model = eng._batch, eng._model
    T = 8
    emb = torch.empty(T, 64)
    RaggedEmbeddingKernel(torch.float32, torch.int64, 64)(emb, batch, model.embed)
    assert torch.equal(emb, sd["model.embed_tokens.weight"][torch.cat([a, b])])
    last = torch.empty(2, 64)
    RaggedLogitsGather(64, torch.float32)(last, emb, batch)
    assert torch.equal(last, emb[[4, 7]])
    atoms, n = AtomBuilder()(None, batch, 128, 16)
    assert n == 2
    # KV copy without rotation, then attention == per-sequence causal softmax(q k^T) v
    hq, hkv, d = 4, 2, 16
    qkv = torch.randn(T, (hq + 2 * hkv) * d)
    cache = torch.zeros(eng._state_manager.get_cache(0).shape)  # fp32 cache: exact comparison
    LinearBlockedKVCopy(d, hq, hkv, torch.float32)(cache, qkv, batch)
    out = torch.empty(T, hq * d)
    BlockedFlashAttn(d, torch.float32)(out, qkv, cache, batch, hq, hkv)
    q, k, v = qkv.view(T, hq + 2 * hkv, d).split([hq, hkv, hkv], 1)
    for lo, hi in ((0, 5), (5, 8)):
        ref = torch.nn.functional.scaled_dot_product_attention(q[lo:hi].transpose(0, 1), k[lo:hi].transpose(0, 1).repeat_interleave(2, 0),
                                                               v[lo:hi].transpose(0, 1).repeat_interleave(2, 0), is_causal=True)
        assert torch.allclose(out[lo:hi], ref.transpose(0, 1).reshape(hi - lo, -1), atol=1e-5)
    rot = qkv.clone()
    BlockedRotaryEmbeddings(d, hq, hkv, torch.float32, d, 10000.0)(cache, rot, batch)
    assert torch.equal(rot[:, (hq + hkv) * d:], qkv[:, (hq + hkv) * d:]) and not torch.equal(rot[1, :d], qkv[1, :d])
    assert torch.allclose(rot[0, :hq * d], qkv[0, :hq * d], atol=1e-6)  # position 0 is not rotated
    # gating -> scatter -> gather with identity experts gives back the input (normalised top-2 weights sum to one)
    E, k2 = 4, 2
    logits = torch.randn(T, E)
    counts, scores, asg, off = torch.empty(E, dtype=torch.int32), torch.empty(T, k2), torch.empty(T, k2, dtype=torch.int32), torch.empty(
        T, k2, dtype=torch.int32)
    RaggedTopKGating(torch.float32)(counts, scores, asg, off, logits, batch)
    assert counts.sum() == T * k2 and torch.allclose(scores.sum(-1), torch.ones(T))
    moe_in, cumsum, slots = torch.empty(T * k2, 64), torch.empty(E, dtype=torch.int64), torch.empty(T, k2, dtype=torch.int32)
    MoEScatter(torch.float32, 64)(moe_in, cumsum, slots, emb, counts, asg, off)
    assert cumsum[-1] == T * k2
    back = torch.empty(T, 64)
    MoEGather(torch.float32, 64)(back, moe_in, scores, slots)
    assert torch.allclose(back, emb, atol=1e-6)
