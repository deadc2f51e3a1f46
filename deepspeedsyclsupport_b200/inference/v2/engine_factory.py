"""Engine construction: HF checkpoint -> canonical parameter names -> RaggedTransformer -> InferenceEngineV2.
Parity: inference/v2/engine_factory.py:28-123 (build_hf_engine, build_engine_from_ds_checkpoint) and the per-family
policies/containers that map HF tensor names (model_implementations/*/policy.py, *_containers.py)."""
import json
import os
import pickle
import re

import torch

from .config_v2 import RaggedInferenceEngineConfig
from .engine_v2 import InferenceEngineV2
from .model_implementations import SPEC_FROM_HF, ArchSpec, RaggedTransformer


def _map_llama(sd, cfg):
    out = {"embed.weight": sd["model.embed_tokens.weight"], "final_norm.weight": sd["model.norm.weight"]}
    if "lm_head.weight" in sd:
        out["lm_head.weight"] = sd["lm_head.weight"]
    for i in range(cfg["num_hidden_layers"]):
        p, q = f"model.layers.{i}.", f"layers.{i}."
        for a, b in (("self_attn.q_proj", "q"), ("self_attn.k_proj", "k"), ("self_attn.v_proj", "v"), ("self_attn.o_proj", "o")):
            out[q + b + ".weight"] = sd[p + a + ".weight"]
            if p + a + ".bias" in sd:  # internlm / llama with attention_bias
                out[q + b + ".bias"] = sd[p + a + ".bias"]
        out[q + "norm1.weight"] = sd[p + "input_layernorm.weight"]
        out[q + "norm2.weight"] = sd[p + "post_attention_layernorm.weight"]
        if p + "mlp.gate_proj.weight" in sd:
            out[q + "gate.weight"], out[q + "up.weight"], out[q + "down.weight"] = (sd[p + "mlp.gate_proj.weight"],
                                                                                     sd[p + "mlp.up_proj.weight"], sd[p + "mlp.down_proj.weight"])
        else:  # mixtral
            out[q + "router.weight"] = sd[p + "block_sparse_moe.gate.weight"]
            for e in range(cfg["num_local_experts"]):
                ep = f"{p}block_sparse_moe.experts.{e}."
                out[f"{q}experts.{e}.gate.weight"] = sd[ep + "w1.weight"]
                out[f"{q}experts.{e}.down.weight"] = sd[ep + "w2.weight"]
                out[f"{q}experts.{e}.up.weight"] = sd[ep + "w3.weight"]
    return out


def _map_opt(sd, cfg):
    pre = "model.decoder." if "model.decoder.embed_tokens.weight" in sd else "decoder."
    out = {"embed.weight": sd[pre + "embed_tokens.weight"], "pos_embed.weight": sd[pre + "embed_positions.weight"],
           "final_norm.weight": sd[pre + "final_layer_norm.weight"], "final_norm.bias": sd[pre + "final_layer_norm.bias"]}
    for i in range(cfg["num_hidden_layers"]):
        p, q = f"{pre}layers.{i}.", f"layers.{i}."
        for a, b in (("self_attn.q_proj", "q"), ("self_attn.k_proj", "k"), ("self_attn.v_proj", "v"), ("self_attn.out_proj", "o"),
                     ("fc1", "up"), ("fc2", "down")):
            out[q + b + ".weight"], out[q + b + ".bias"] = sd[p + a + ".weight"], sd[p + a + ".bias"]
        out[q + "norm1.weight"], out[q + "norm1.bias"] = sd[p + "self_attn_layer_norm.weight"], sd[p + "self_attn_layer_norm.bias"]
        out[q + "norm2.weight"], out[q + "norm2.bias"] = sd[p + "final_layer_norm.weight"], sd[p + "final_layer_norm.bias"]
    return out


def _map_falcon(sd, cfg):
    hq = cfg["num_attention_heads"]
    hkv = cfg.get("num_kv_heads", 1 if cfg.get("multi_query", True) else hq)
    d = cfg["hidden_size"] // hq
    out = {"embed.weight": sd["transformer.word_embeddings.weight"], "final_norm.weight": sd["transformer.ln_f.weight"],
           "final_norm.bias": sd["transformer.ln_f.bias"]}
    if "lm_head.weight" in sd:
        out["lm_head.weight"] = sd["lm_head.weight"]
    for i in range(cfg["num_hidden_layers"]):
        p, q = f"transformer.h.{i}.", f"layers.{i}."
        w = sd[p + "self_attention.query_key_value.weight"]
        if hkv == hq:  # interleaved [h, 3, d]
            w3 = w.view(hq, 3, d, -1)
            out[q + "q.weight"], out[q + "k.weight"], out[q + "v.weight"] = (w3[:, 0].reshape(hq * d, -1), w3[:, 1].reshape(hq * d, -1),
                                                                             w3[:, 2].reshape(hq * d, -1))
        else:  # grouped: [hkv, group + 2, d]
            g = hq // hkv
            wg = w.view(hkv, g + 2, d, -1)
            out[q + "q.weight"] = wg[:, :g].reshape(hq * d, -1)
            out[q + "k.weight"] = wg[:, g].reshape(hkv * d, -1)
            out[q + "v.weight"] = wg[:, g + 1].reshape(hkv * d, -1)
        out[q + "o.weight"] = sd[p + "self_attention.dense.weight"]
        ln = "input_layernorm" if p + "input_layernorm.weight" in sd else "ln_attn"
        out[q + "norm1.weight"], out[q + "norm1.bias"] = sd[p + ln + ".weight"], sd[p + ln + ".bias"]
        out[q + "up.weight"], out[q + "down.weight"] = sd[p + "mlp.dense_h_to_4h.weight"], sd[p + "mlp.dense_4h_to_h.weight"]
    return out


def _map_phi(sd, cfg):
    out = {"embed.weight": sd["model.embed_tokens.weight"], "final_norm.weight": sd["model.final_layernorm.weight"],
           "final_norm.bias": sd["model.final_layernorm.bias"], "lm_head.weight": sd["lm_head.weight"]}
    for i in range(cfg["num_hidden_layers"]):
        p, q = f"model.layers.{i}.", f"layers.{i}."
        for a, b in (("self_attn.q_proj", "q"), ("self_attn.k_proj", "k"), ("self_attn.v_proj", "v"), ("self_attn.dense", "o"),
                     ("mlp.fc1", "up"), ("mlp.fc2", "down")):
            out[q + b + ".weight"], out[q + b + ".bias"] = sd[p + a + ".weight"], sd[p + a + ".bias"]
        out[q + "norm1.weight"], out[q + "norm1.bias"] = sd[p + "input_layernorm.weight"], sd[p + "input_layernorm.bias"]
    return out


def _map_gpt2(sd, cfg):
    """GPT-2: Conv1D weights are stored [in, out]; c_attn is q | k | v along the output axis."""
    pre = "transformer." if "transformer.wte.weight" in sd else ""
    h = cfg["n_embd"]
    out = {"embed.weight": sd[pre + "wte.weight"], "pos_embed.weight": sd[pre + "wpe.weight"], "final_norm.weight": sd[pre + "ln_f.weight"],
           "final_norm.bias": sd[pre + "ln_f.bias"]}
    for i in range(cfg["n_layer"]):
        p, q = f"{pre}h.{i}.", f"layers.{i}."
        w, b = sd[p + "attn.c_attn.weight"].t(), sd[p + "attn.c_attn.bias"]
        for j, n in enumerate("qkv"):
            out[q + n + ".weight"], out[q + n + ".bias"] = w[j * h:(j + 1) * h].contiguous(), b[j * h:(j + 1) * h]
        out[q + "o.weight"], out[q + "o.bias"] = sd[p + "attn.c_proj.weight"].t().contiguous(), sd[p + "attn.c_proj.bias"]
        out[q + "up.weight"], out[q + "up.bias"] = sd[p + "mlp.c_fc.weight"].t().contiguous(), sd[p + "mlp.c_fc.bias"]
        out[q + "down.weight"], out[q + "down.bias"] = sd[p + "mlp.c_proj.weight"].t().contiguous(), sd[p + "mlp.c_proj.bias"]
        for a, b2 in (("ln_1", "norm1"), ("ln_2", "norm2")):
            out[q + b2 + ".weight"], out[q + b2 + ".bias"] = sd[p + a + ".weight"], sd[p + a + ".bias"]
    return out


def _map_gptj(sd, cfg):
    """GPT-J rotates interleaved pairs (2i, 2i+1) of the first rotary_dim features of every head; the kernels rotate
    (i, i + rot/2).  Permuting those rows of W_q and W_k the same way (evens first, then odds) leaves q.k unchanged."""
    import torch
    hq = cfg["n_head"]
    d = cfg["n_embd"] // hq
    rot = cfg.get("rotary_dim") or d
    perm = torch.cat([torch.arange(0, rot, 2), torch.arange(1, rot, 2), torch.arange(rot, d)])
    rows = (torch.arange(hq)[:, None] * d + perm[None, :]).reshape(-1)
    out = {"embed.weight": sd["transformer.wte.weight"], "final_norm.weight": sd["transformer.ln_f.weight"],
           "final_norm.bias": sd["transformer.ln_f.bias"], "lm_head.weight": sd["lm_head.weight"]}
    if "lm_head.bias" in sd:
        out["lm_head.bias"] = sd["lm_head.bias"]
    for i in range(cfg["n_layer"]):
        p, q = f"transformer.h.{i}.", f"layers.{i}."
        out[q + "q.weight"], out[q + "k.weight"] = sd[p + "attn.q_proj.weight"][rows], sd[p + "attn.k_proj.weight"][rows]
        out[q + "v.weight"], out[q + "o.weight"] = sd[p + "attn.v_proj.weight"], sd[p + "attn.out_proj.weight"]
        out[q + "up.weight"], out[q + "up.bias"] = sd[p + "mlp.fc_in.weight"], sd[p + "mlp.fc_in.bias"]
        out[q + "down.weight"], out[q + "down.bias"] = sd[p + "mlp.fc_out.weight"], sd[p + "mlp.fc_out.bias"]
        out[q + "norm1.weight"], out[q + "norm1.bias"] = sd[p + "ln_1.weight"], sd[p + "ln_1.bias"]
    return out


def _map_gpt_neox(sd, cfg):
    hq = cfg["num_attention_heads"]
    d = cfg["hidden_size"] // hq
    out = {"embed.weight": sd["gpt_neox.embed_in.weight"], "final_norm.weight": sd["gpt_neox.final_layer_norm.weight"],
           "final_norm.bias": sd["gpt_neox.final_layer_norm.bias"], "lm_head.weight": sd["embed_out.weight"]}
    for i in range(cfg["num_hidden_layers"]):
        p, q = f"gpt_neox.layers.{i}.", f"layers.{i}."
        w = sd[p + "attention.query_key_value.weight"].view(hq, 3, d, -1)  # interleaved [head, (q k v), d]
        for j, n in enumerate("qkv"):
            out[q + n + ".weight"] = w[:, j].reshape(hq * d, -1)
        if p + "attention.query_key_value.bias" in sd:
            b = sd[p + "attention.query_key_value.bias"].view(hq, 3, d)
            for j, n in enumerate("qkv"):
                out[q + n + ".bias"] = b[:, j].reshape(hq * d)
        for a, n in (("attention.dense", "o"), ("mlp.dense_h_to_4h", "up"), ("mlp.dense_4h_to_h", "down")):
            out[q + n + ".weight"] = sd[p + a + ".weight"]
            if p + a + ".bias" in sd:
                out[q + n + ".bias"] = sd[p + a + ".bias"]
        for a, n in (("input_layernorm", "norm1"), ("post_attention_layernorm", "norm2")):
            out[q + n + ".weight"], out[q + n + ".bias"] = sd[p + a + ".weight"], sd[p + a + ".bias"]
    return out


_MAPPERS = {"llama": _map_llama, "mistral": _map_llama, "internlm": _map_llama, "mixtral": _map_llama, "opt": _map_opt,
            "falcon": _map_falcon, "phi": _map_phi, "gpt2": _map_gpt2, "gptj": _map_gptj, "gpt_neox": _map_gpt_neox}


def build_engine_from_state_dict(model_type, hf_config: dict, state_dict: dict, engine_config: RaggedInferenceEngineConfig,
                                 dtype=torch.bfloat16):
    """Engine from an in-memory HF-named state dict (used for random-init models when no checkpoint is on disk)."""
    if model_type not in SPEC_FROM_HF:
        raise ValueError(f"Unsupported model type {model_type}")
    spec = SPEC_FROM_HF[model_type](hf_config)
    canon = _MAPPERS[model_type](state_dict, hf_config)

    def builder(tp_group, device):
        return RaggedTransformer(spec, canon, tp_group=tp_group, dtype=dtype, device=device,
                                 quantization_mode=engine_config.quantization.quantization_mode, engine_config=engine_config)

    return InferenceEngineV2(builder, engine_config)


def build_hf_engine(path: str, engine_config: RaggedInferenceEngineConfig, debug_level: int = 20):
    if os.path.exists(os.path.join(path, "ds_model_config.pkl")):
        return build_engine_from_ds_checkpoint(path, engine_config, debug_level=debug_level)
    from .checkpoint.huggingface_engine import HuggingFaceCheckpointEngine
    ck = HuggingFaceCheckpointEngine(path)
    model_type = ck.model_config["model_type"]
    sd = dict(ck.parameters())
    return build_engine_from_state_dict(model_type, ck.model_config, sd, engine_config)


def build_engine_from_ds_checkpoint(path: str, engine_config: RaggedInferenceEngineConfig, debug_level: int = 20):
    """Rebuild from `InferenceEngineV2.serialize()` output: per rank one flat byte buffer (already TP-sharded and laid out for the
    kernels, fp8 scales included) + its description; the containers are pointed at slices of the buffer."""
    from .inference_parameter import STR_TO_DTYPE
    from .model_implementations.flat_model_helpers import (make_metadata_filename, make_model_config_filename, make_param_filename,
                                                           restore_inference_model)
    with open(make_model_config_filename(path), "rb") as f:
        saved = pickle.load(f)

    def builder(tp_group, device):
        m = RaggedTransformer(saved["spec"], None, tp_group=tp_group, dtype=STR_TO_DTYPE[saved["dtype"]], device=device,
                              quantization_mode=saved.get("quantization_mode", ""), engine_config=engine_config)
        buffer = torch.load(make_param_filename(path, m.tp_rank, m.tp_size), map_location="cpu").to(device)
        with open(make_metadata_filename(path, m.tp_rank, m.tp_size)) as f:
            restore_inference_model(buffer, json.load(f), m.named_containers())
        m.flattened_params = buffer
        m.check_populated()
        return m

    return InferenceEngineV2(builder, engine_config)
