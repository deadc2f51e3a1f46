"""FastGen v2 engine on the host path: ragged put() with prefill + decode + chunked prefill must reproduce a dense
causal forward of the same random-init Llama weights (HF naming)."""
import pytest
import torch


def _hf_llama_state(cfg, seed=0):
    g = torch.Generator().manual_seed(seed)
    h, inter, L, V = cfg["hidden_size"], cfg["intermediate_size"], cfg["num_hidden_layers"], cfg["vocab_size"]
    hq, hkv = cfg["num_attention_heads"], cfg["num_key_value_heads"]
    d = h // hq
    r = lambda *s: torch.randn(*s, generator=g) * 0.05
    sd = {"model.embed_tokens.weight": r(V, h), "model.norm.weight": 1 + r(h), "lm_head.weight": r(V, h)}
    for i in range(L):
        p = f"model.layers.{i}."
        sd[p + "self_attn.q_proj.weight"] = r(hq * d, h)
        sd[p + "self_attn.k_proj.weight"] = r(hkv * d, h)
        sd[p + "self_attn.v_proj.weight"] = r(hkv * d, h)
        sd[p + "self_attn.o_proj.weight"] = r(h, hq * d)
        sd[p + "input_layernorm.weight"] = 1 + r(h)
        sd[p + "post_attention_layernorm.weight"] = 1 + r(h)
        sd[p + "mlp.gate_proj.weight"] = r(inter, h)
        sd[p + "mlp.up_proj.weight"] = r(inter, h)
        sd[p + "mlp.down_proj.weight"] = r(h, inter)
    return sd


def _dense_llama_logits(sd, cfg, ids):
    """fp32 dense causal forward -> logits for every position [S, V]."""
    import torch.nn.functional as F
    from deepspeedsyclsupport_b200.ops.fused_ops import rope_inv_freq, rope_ref
    h, hq, hkv = cfg["hidden_size"], cfg["num_attention_heads"], cfg["num_key_value_heads"]
    d = h // hq
    S = ids.numel()
    x = sd["model.embed_tokens.weight"][ids].float()
    inv = rope_inv_freq(d)
    pos = torch.arange(S)
    rms = lambda t, w: t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + 1e-5) * w
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        hn = rms(x, sd[p + "input_layernorm.weight"])
        qkv = torch.cat([hn @ sd[p + "self_attn.q_proj.weight"].t(), hn @ sd[p + "self_attn.k_proj.weight"].t(),
                         hn @ sd[p + "self_attn.v_proj.weight"].t()], -1)
        qkv = rope_ref(qkv, inv, pos, 0, hq + hkv, d)
        q, k, v = qkv.view(S, hq + 2 * hkv, d).split([hq, hkv, hkv], 1)
        o = F.scaled_dot_product_attention(q.transpose(0, 1)[None], k.transpose(0, 1).repeat_interleave(hq // hkv, 0)[None],
                                           v.transpose(0, 1).repeat_interleave(hq // hkv, 0)[None], is_causal=True)[0]
        x = x + o.transpose(0, 1).reshape(S, -1) @ sd[p + "self_attn.o_proj.weight"].t()
        hn = rms(x, sd[p + "post_attention_layernorm.weight"])
        x = x + (F.silu(hn @ sd[p + "mlp.gate_proj.weight"].t()) * (hn @ sd[p + "mlp.up_proj.weight"].t())) @ sd[p + "mlp.down_proj.weight"].t()
    return rms(x, sd["model.norm.weight"]) @ sd["lm_head.weight"].t()


CFG = dict(vocab_size=97, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
           rms_norm_eps=1e-5)


def _engine(dtype=torch.float32):
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64, "max_ragged_sequence_count": 8, "max_context": 128,
                                                      "memory_config": {"mode": "allocate", "size": 40}})
    sd = _hf_llama_state(CFG)
    return build_engine_from_state_dict("llama", CFG, sd, ecfg, dtype=dtype), sd


def test_ragged_prefill_decode_matches_dense():
    eng, sd = _engine()
    torch.manual_seed(1)
    a, b = torch.randint(0, 97, (21, )), torch.randint(0, 97, (9, ))
    ref_a, ref_b = _dense_llama_logits(sd, CFG, a), _dense_llama_logits(sd, CFG, b)
    # step 1: prefill a[:17] (chunk), full prefill of b minus its last token
    out = eng.put([1, 2], [a[:17], b[:8]])
    assert torch.allclose(out[0], ref_a[16], atol=5e-3) and torch.allclose(out[1], ref_b[7], atol=5e-3)
    # step 2: rest of a's prompt chunk fused with a decode token of b
    out = eng.put([1, 2], [a[17:20], b[8:9]])
    assert torch.allclose(out[0], ref_a[19], atol=5e-3) and torch.allclose(out[1], ref_b[8], atol=5e-3)
    out = eng.put([1], [a[20:21]])
    assert torch.allclose(out[0], ref_a[20], atol=5e-3)
    free_before = eng.free_blocks
    eng.flush(2)
    assert eng.free_blocks > free_before
    assert eng.query(1, 5, 10)[0] == 5


def test_scheduling_limits_and_splitfuse():
    from deepspeedsyclsupport_b200.inference.v2 import SchedulingResult, SplitFuseScheduler
    eng, sd = _engine()
    assert eng.can_schedule([1], [65]) == SchedulingResult.BatchTokenLimitExceeded
    assert eng.can_schedule(list(range(9)), [1] * 9) == SchedulingResult.BatchSequenceLimitExceeded
    assert eng.can_schedule([1], [200]) == SchedulingResult.KVCacheLimitExceeded
    sch = SplitFuseScheduler(eng, token_budget=16)
    torch.manual_seed(3)
    prompts = {10: torch.randint(0, 97, (40, )), 11: torch.randint(0, 97, (5, ))}
    for uid, p in prompts.items():
        sch.add_request(uid, p, max_new_tokens=3)
    done = sch.run_to_completion()
    for uid, p in prompts.items():
        seq = p.clone()
        for t in done[uid]:
            nxt = int(_dense_llama_logits(sd, CFG, seq)[-1].argmax())
            assert nxt == t
            seq = torch.cat([seq, torch.tensor([t])])
    assert eng.free_blocks == 40


MOE_CFG = dict(CFG, num_local_experts=4, num_experts_per_tok=2)


def _hf_mixtral_state(cfg, seed=0):
    sd = _hf_llama_state(cfg, seed)
    g = torch.Generator().manual_seed(seed + 100)
    h, inter = cfg["hidden_size"], cfg["intermediate_size"]
    r = lambda *s: torch.randn(*s, generator=g) * 0.05
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        for k in ("gate_proj", "up_proj", "down_proj"):
            del sd[p + f"mlp.{k}.weight"]
        sd[p + "block_sparse_moe.gate.weight"] = r(cfg["num_local_experts"], h) * 10
        for e in range(cfg["num_local_experts"]):
            ep = f"{p}block_sparse_moe.experts.{e}."
            sd[ep + "w1.weight"], sd[ep + "w2.weight"], sd[ep + "w3.weight"] = r(inter, h), r(h, inter), r(inter, h)
    return sd


def _dense_mixtral_logits(sd, cfg, ids):
    """the Llama reference with the MLP swapped for a dense evaluation of the top-2 mixture"""
    import torch.nn.functional as F
    E, k = cfg["num_local_experts"], cfg["num_experts_per_tok"]
    from deepspeedsyclsupport_b200.ops.fused_ops import rope_inv_freq, rope_ref
    h, hq, hkv = cfg["hidden_size"], cfg["num_attention_heads"], cfg["num_key_value_heads"]
    d = h // hq
    S = ids.numel()
    x = sd["model.embed_tokens.weight"][ids].float()
    inv, pos = rope_inv_freq(d), torch.arange(S)
    rms = lambda t, w: t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + 1e-5) * w
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        hn = rms(x, sd[p + "input_layernorm.weight"])
        qkv = torch.cat([hn @ sd[p + "self_attn.q_proj.weight"].t(), hn @ sd[p + "self_attn.k_proj.weight"].t(),
                         hn @ sd[p + "self_attn.v_proj.weight"].t()], -1)
        qkv = rope_ref(qkv, inv, pos, 0, hq + hkv, d)
        q, kk, v = qkv.view(S, hq + 2 * hkv, d).split([hq, hkv, hkv], 1)
        o = F.scaled_dot_product_attention(q.transpose(0, 1)[None], kk.transpose(0, 1).repeat_interleave(hq // hkv, 0)[None],
                                           v.transpose(0, 1).repeat_interleave(hq // hkv, 0)[None], is_causal=True)[0]
        x = x + o.transpose(0, 1).reshape(S, -1) @ sd[p + "self_attn.o_proj.weight"].t()
        hn = rms(x, sd[p + "post_attention_layernorm.weight"])
        probs = F.softmax(hn @ sd[p + "block_sparse_moe.gate.weight"].t(), -1)
        top, idx = probs.topk(k, -1)
        top = top / top.sum(-1, keepdim=True)
        y = torch.zeros_like(hn)
        for e in range(E):
            ep = f"{p}block_sparse_moe.experts.{e}."
            ye = (F.silu(hn @ sd[ep + "w1.weight"].t()) * (hn @ sd[ep + "w3.weight"].t())) @ sd[ep + "w2.weight"].t()
            wgt = (top * (idx == e)).sum(-1, keepdim=True)
            y = y + wgt * ye
        x = x + y
    return rms(x, sd["model.norm.weight"]) @ sd["lm_head.weight"].t()


def test_ragged_mixtral_grouped_experts_match_dense():
    """top-2 MoE through gating -> scatter -> grouped expert GEMM -> gather (reference: model_implementations/mixtral)."""
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64, "max_ragged_sequence_count": 8, "max_context": 128,
                                                      "memory_config": {"mode": "allocate", "size": 40}})
    sd = _hf_mixtral_state(MOE_CFG)
    eng = build_engine_from_state_dict("mixtral", MOE_CFG, sd, ecfg, dtype=torch.float32)
    torch.manual_seed(4)
    a, b = torch.randint(0, 97, (19, )), torch.randint(0, 97, (6, ))
    ref_a, ref_b = _dense_mixtral_logits(sd, MOE_CFG, a), _dense_mixtral_logits(sd, MOE_CFG, b)
    out = eng.put([1, 2], [a[:18], b])
    assert torch.allclose(out[0], ref_a[17], atol=5e-3) and torch.allclose(out[1], ref_b[5], atol=5e-3)
    out = eng.put([1], [a[18:19]])
    assert torch.allclose(out[0], ref_a[18], atol=5e-3)


def test_ragged_surface_accessors_and_kv_offload():
    """reference-named accessors of the ragged batch / sequence descriptor / KV cache (offload -> restore round trip)."""
    eng, _ = _engine()
    sm = eng._state_manager
    torch.manual_seed(0)
    eng.put([7], [torch.randint(0, 97, (20, ))])
    seq = sm.get_sequence(7)
    assert seq.tracking_id == 7 and seq.max_context >= 128 and seq.kv_cache_ids().numel() == seq.cur_allocated_blocks
    assert sm.n_kv_cache_groups == 1 and sm._kv_cache.num_caches == 1
    ids = seq.kv_cache_ids().tolist()
    before = sm.get_cache(0)[ids].clone()
    host = sm._kv_cache.offload(ids)
    sm.get_cache(0)[ids] = 0
    sm.get_cache(1)[ids] = 0
    sm._kv_cache.restore(ids, host)
    assert torch.equal(sm.get_cache(0)[ids], before)
    b = eng._batch
    b.clear()
    b.insert_sequence(seq, torch.randint(0, 97, (3, )))
    b.finalize()
    assert b.tensor_toks == 3 and b.tokens_to_seq(False).tolist() == [0, 0, 0]
    assert b.batch_metadata_buffer(False).tolist() == [3, 1] and b.inflight_seq_descriptors(False)[0].tolist() == [0, 3, 20, 0]
    assert b.kv_ptrs(False).shape[0] == 1 and b.masks() is None
    n = seq.cur_allocated_blocks
    seq.free_kv_cache(ids[:1])
    assert seq.cur_allocated_blocks == n - 1 and seq.kv_cache_ids().tolist() == ids[1:]


def test_ragged_engine_w8a8_mode_tracks_dense():
    """quantization_mode = wf8af8: projections run as fp8 x fp8 with row scales (host restatement here, tcgen05 kernel on the
    device); logits stay within fp8 tolerance of the dense fp32 model and greedy tokens mostly agree."""
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64, "max_ragged_sequence_count": 8, "max_context": 128,
                                                      "memory_config": {"mode": "allocate", "size": 40}},
                                       quantization={"quantization_mode": "wf8af8"})
    sd = _hf_llama_state(CFG)
    eng = build_engine_from_state_dict("llama", CFG, sd, ecfg, dtype=torch.float32)
    qkv_w = eng._model.layers[0]["qkv_w"]
    assert qkv_w.dtype == torch.float8_e4m3fn and qkv_w.scales.shape == (qkv_w.shape[0], )
    assert type(eng._model.qkv_lin).__name__ == "FP8Linear"  # picked by modules/heuristics from the engine's quantization mode
    torch.manual_seed(1)
    a = torch.randint(0, 97, (21, ))
    ref = _dense_llama_logits(sd, CFG, a)
    out = eng.put([1], [a])
    assert (out[0].float() - ref[20]).abs().max() < 0.2 * ref[20].abs().max()
