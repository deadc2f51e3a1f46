"""End-to-end workloads for the BASELINE.json configs that are not the headline (bench.py runs the headline, config #2):

    --workload fastgen70b   config #5  Llama-2-70B shapes, FastGen ragged batching, tensor parallel over all ranks (TP = world),
                                       bf16 or block-scaled fp8 (--quant wf8af8_mx | wf8af8_block | wf8af8), synthetic ragged arrivals
    --workload mixtral      config #4  Mixtral 8x7B shapes, expert parallel over all ranks, ZeRO-2 training step
    --workload ulysses13b   config #3  Llama-2-13B shapes, Ulysses sequence parallel over all ranks, ZeRO-3, --seq-len tokens total

Launched like bench.py (`python -m torch.distributed.run ... bench.py --workload X --gpus N ...`).  Random-init weights, synthetic
tokens.  Every timed region is end to end: the step's inputs start in pinned host memory, the step's result (sampled tokens /
loss) is read back to the host; timing = CUDA events on the device, max over ranks, barrier + synchronize on both sides.  Rank 0
prints one JSON line in the bench.py vocabulary.  `--layers L` cuts the depth for shake-down runs (the model name then says so)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def _timed(torch, dist, fn, steps, warmup):
    """-> (ms per call of fn, max over ranks, device-timed), clocks dict (rank 0)."""
    from bench import ClockSampler
    dev = torch.device("cuda", torch.cuda.current_device())
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    sampler = ClockSampler(torch.cuda.current_device())
    if dist.get_rank() == 0:
        sampler.start()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([a.elapsed_time(b)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if dist.get_rank() == 0 else None
    return ms.item() / steps, clocks


def _line(metric, value, unit, world, args, ms, config, clocks, launches, extra):
    out = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
           "data": "synthetic tokens (uniform random ids), random-init weights", "impl": "own", "config": config,
           "gpu_launches": launches, "clocks": clocks}
    out.update(extra)
    return out


# ------------------------------------------------------------------------------------------------ config #5: FastGen 70B
class _LazyLlamaState(dict):
    """HF-named Llama state dict whose tensors are drawn on the device when they are asked for (a 70B checkpoint is 140 GB; the
    engine's loader shards and drops every tensor as soon as its parameter is complete)."""

    def __init__(self, cfg, device, dtype):
        super().__init__()
        self.cfg, self.device, self.dtype = cfg, device, dtype

    def items(self):
        import torch
        c = self.cfg
        h, inter, V = c["hidden_size"], c["intermediate_size"], c["vocab_size"]
        d = h // c["num_attention_heads"]
        hq, hkv = c["num_attention_heads"], c["num_key_value_heads"]

        def r(*shape, std=0.02, one=False):
            t = torch.empty(*shape, device=self.device, dtype=self.dtype).normal_(0.0, std)
            return t + 1 if one else t

        yield "model.embed_tokens.weight", r(V, h)
        for i in range(c["num_hidden_layers"]):
            p = f"model.layers.{i}."
            yield p + "self_attn.q_proj.weight", r(hq * d, h)
            yield p + "self_attn.k_proj.weight", r(hkv * d, h)
            yield p + "self_attn.v_proj.weight", r(hkv * d, h)
            yield p + "self_attn.o_proj.weight", r(h, hq * d)
            yield p + "input_layernorm.weight", r(h, one=True)
            yield p + "post_attention_layernorm.weight", r(h, one=True)
            yield p + "mlp.gate_proj.weight", r(inter, h)
            yield p + "mlp.up_proj.weight", r(inter, h)
            yield p + "mlp.down_proj.weight", r(h, inter)
        yield "model.norm.weight", r(h, one=True)
        yield "lm_head.weight", r(V, h)


def run_fastgen70b(args):
    sys.path.insert(0, ROOT)
    import torch
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    from deepspeedsyclsupport_b200.ops import native
    deepspeed.init_distributed()
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    L = args.layers or 80
    cfg = dict(vocab_size=32000, hidden_size=8192, intermediate_size=28672, num_hidden_layers=L, num_attention_heads=64,
               num_key_value_heads=8, rms_norm_eps=1e-5, rope_theta=10000.0, max_position_embeddings=4096)
    n_seqs, budget, gen = args.sequences, args.token_budget, args.gen_tokens
    g = torch.Generator().manual_seed(7)
    prompt_lens = torch.randint(args.prompt_min, args.prompt_max + 1, (n_seqs, ), generator=g).tolist()
    max_ctx = max(prompt_lens) + gen * (args.steps + args.warmup) + 8
    blocks = n_seqs * (-(-max_ctx // 128) + 1) + 8
    ecfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": world},
                                       state_manager={"max_ragged_batch_size": budget, "max_ragged_sequence_count": max(n_seqs, 8),
                                                      "max_tracked_sequences": max(n_seqs, 8), "max_context": -(-max_ctx // 128) * 128,
                                                      "memory_config": {"mode": "allocate", "size": blocks}},
                                       quantization={"quantization_mode": args.quant})
    t_build = time.perf_counter()
    eng = build_engine_from_state_dict("llama", cfg, _LazyLlamaState(cfg, dev, torch.bfloat16), ecfg, dtype=torch.bfloat16)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build
    prompts = [torch.randint(0, cfg["vocab_size"], (n, ), generator=g).pin_memory() for n in prompt_lens]
    uids = list(range(n_seqs))
    state = {"next": None, "first": True}

    def prefill():
        """Ragged arrivals: every forward takes whole prompts while they fit the token budget, then the next prompt's head as a
        chunk (split-fuse); the unfinished prompt continues in the next forward."""
        for u in uids:
            if not state["first"]:
                eng.flush(u)
        state["first"] = False
        nxt = torch.zeros(n_seqs, dtype=torch.long)
        pos = [0] * n_seqs
        i = 0
        forwards = 0
        while i < n_seqs:
            batch_u, batch_t, room = [], [], budget
            j = i
            while j < n_seqs and room > 0 and len(batch_u) < ecfg.state_manager.max_ragged_sequence_count:
                take = min(room, prompt_lens[j] - pos[j])
                batch_u.append(j)
                batch_t.append(prompts[j][pos[j]:pos[j] + take])
                pos[j] += take
                room -= take
                if pos[j] < prompt_lens[j]:
                    break
                j += 1
            logits = eng.put(batch_u, batch_t)
            tok = logits.argmax(-1).cpu()  # device -> host: the sampled tokens
            for k, u in enumerate(batch_u):
                if pos[u] == prompt_lens[u]:
                    nxt[u] = tok[k]
            forwards += 1
            i = j
        state["next"] = nxt
        return forwards

    def decode():
        nxt = state["next"]
        for _ in range(gen):
            logits = eng.put(uids, [nxt[u:u + 1] for u in uids])
            nxt = logits.argmax(-1).cpu()
        state["next"] = nxt

    l0 = native.launch_count
    ms_prefill, _ = _timed(torch, dist, prefill, 1, 1)  # one untimed pass (allocations, tensor maps), one timed
    ms_decode, clocks = _timed(torch, dist, decode, args.steps, args.warmup)
    launches = native.launch_count - l0
    if rank == 0:
        prompt_tokens = sum(prompt_lens)
        ctx = sum(prompt_lens) / n_seqs + gen * (args.warmup + args.steps / 2)
        ops = eng._model._tp_ops() if hasattr(eng._model, "_tp_ops") else None
        extra = {
            "prefill": {"prompt_tokens": prompt_tokens, "ms": round(ms_prefill, 2), "tokens_per_s": round(prompt_tokens / ms_prefill * 1e3, 1)},
            "decode": {"ms_per_token_step": round(ms_decode / gen, 3), "sequences": n_seqs, "mean_context": round(ctx, 1),
                       "tokens_per_s": round(n_seqs * gen / ms_decode * 1e3, 1)},
            "e2e": {"value": round(n_seqs * gen / ms_decode * 1e3, 1), "unit": "generated tokens/s",
                    "h2d_bytes_per_step": n_seqs * 8, "d2h_bytes_per_step": n_seqs * 8},
            "build_s": round(t_build, 1), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
            "tp_collectives": None if ops is None else dict(ops.calls, nvls=bool(ops.nvls)),
            "reference_arm": "not run: the reference's FastGen (inference v2) needs its CUTLASS / flash-attn extension builds and a 140 GB HF "
                             "checkpoint; its v1 AutoTP path needs the 70B HF module materialised per rank",
        }
        config = {"model": "llama-2-70b" + (f"-{L}layers-INVALID-debug" if args.layers else ""), "parallelism": f"tp{world}",
                  "quantization": args.quant or "bf16", "sequences": n_seqs, "prompt_len": [args.prompt_min, args.prompt_max],
                  "token_budget_per_forward": budget, "generated_tokens_per_timed_step": gen, "kv_block": 128,
                  "l2_policy": "weights (17.5 GB / rank at TP=8) and KV far exceed the 126 MB L2; no explicit flush"}
        print(json.dumps(_line("generated tokens/sec/box Llama-2-70B FastGen ragged decode (TP over all GPUs); prefill tokens/s alongside",
                               n_seqs * gen / ms_decode * 1e3, "tokens/s", world, args, ms_decode, config, clocks, launches, extra)), flush=True)


# ------------------------------------------------------------------------------------------------ config #4: Mixtral EP
def run_mixtral(args):
    sys.path.insert(0, ROOT)
    import torch
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models.mixtral import MixtralForCausalLM, mixtral_config
    from deepspeedsyclsupport_b200.ops import native
    deepspeed.init_distributed()
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    over = {"ep_size": min(world, 8), "activation_checkpointing": True}
    if args.layers:
        over["num_hidden_layers"] = args.layers
    cfg = mixtral_config("8x7b", **over)
    torch.manual_seed(1234 + rank)  # experts differ per rank; the engine broadcasts the shared parameters from rank 0
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = MixtralForCausalLM(cfg)
    torch.set_default_dtype(prev)
    ds = {"train_micro_batch_size_per_gpu": args.micro_batch, "gradient_accumulation_steps": 1, "steps_per_print": 1000000,
          "bf16": {"enabled": True}, "gradient_clipping": 1.0,
          "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
          "zero_optimization": {"stage": 2, "overlap_comm": True, "contiguous_gradients": True, "reduce_bucket_size": 5e8}}
    engine, _, _, _ = deepspeed.initialize(model=model, config=ds)
    B, S, V = args.micro_batch, args.seq_len, cfg.vocab_size
    g = torch.Generator().manual_seed(1000 + rank)
    host_ids = torch.randint(0, V, (B, S), generator=g).pin_memory()
    host_labels = torch.randint(0, V, (B, S), generator=g).pin_memory()
    last = {}

    def step():
        ids, labels = host_ids.to(dev, non_blocking=True), host_labels.to(dev, non_blocking=True)
        loss = engine(ids, labels)
        engine.backward(loss)
        engine.step()
        last["loss"] = loss.item()

    l0 = native.launch_count
    ms, clocks = _timed(torch, dist, step, args.steps, args.warmup)
    launches = native.launch_count - l0
    if rank == 0:
        from deepspeedsyclsupport_b200.moe import sharded_moe as SM
        tokens = world * B * S
        n_params = sum(p.numel() for p in model.parameters())
        extra = {"e2e": {"value": tokens / ms * 1e3, "unit": "tokens/s", "h2d_bytes_per_step": 2 * B * S * 8, "d2h_bytes_per_step": 4},
                 "final_loss": last.get("loss"), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
                 "params_per_rank_b": round(n_params / 1e9, 2),
                 "moe_transport": "peer dispatch-push / combine-pull kernels" if os.environ.get("DSB200_MOE_PEER", "1") != "0" else "nccl all-to-all",
                 "reference_arm": "not run (the reference's MoE + ZeRO-2 on this model needs its fused-kernel JIT builds, which fail for sm_100: see DESIGN.md)"}
        config = {"model": "mixtral-8x7b" + (f"-{args.layers}layers-INVALID-debug" if args.layers else ""), "global_batch": B * world, "seq_len": S,
                  "micro_batch_per_gpu": B, "parallelism": f"ep{min(world, 8)}-zero2-dp{world}", "top_k": 2, "capacity_factor": cfg.capacity_factor,
                  "optimizer": "AdamW", "grad_clip": 1.0, "activation_checkpointing": "full",
                  "l2_policy": "working set far exceeds the 126 MB L2; no explicit flush"}
        line = _line("tokens/sec/box Mixtral-8x7B expert-parallel training step (fwd+bwd+AdamW, ZeRO-2)", tokens / ms * 1e3, "tokens/s", world, args,
                     ms, config, clocks, launches, extra)
        line["scaling"] = "weak"
        print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ config #3: Ulysses 13B
def run_ulysses13b(args):
    sys.path.insert(0, ROOT)
    import torch
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    from deepspeedsyclsupport_b200.ops import native
    deepspeed.init_distributed()
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    over = {"activation_checkpointing": True}
    if args.layers:
        over["num_hidden_layers"] = args.layers
    cfg = llama_config("13b", **over)
    assert cfg.num_attention_heads % world == 0, f"{cfg.num_attention_heads} heads do not split over {world} ranks"
    torch.manual_seed(1234)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg)
    torch.set_default_dtype(prev)
    model.set_sequence_parallel(dist.get_world_group())
    ds = {"train_micro_batch_size_per_gpu": 1, "gradient_accumulation_steps": 1, "steps_per_print": 1000000, "bf16": {"enabled": True},
          "gradient_clipping": 1.0,
          "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
          "zero_optimization": {"stage": 3, "overlap_comm": True, "contiguous_gradients": True, "stage3_prefetch_bucket_size": 5e8,
                                "stage3_max_live_parameters": 1e10, "stage3_param_persistence_threshold": 1e5, "reduce_bucket_size": 5e8}}
    engine, _, _, _ = deepspeed.initialize(model=model, config=ds)
    S = args.seq_len
    S_local = S // world
    g = torch.Generator().manual_seed(1000)
    host_ids = torch.randint(0, cfg.vocab_size, (1, S), generator=g)[:, rank * S_local:(rank + 1) * S_local].contiguous().pin_memory()
    host_labels = torch.randint(0, cfg.vocab_size, (1, S), generator=g)[:, rank * S_local:(rank + 1) * S_local].contiguous().pin_memory()
    last = {}

    def step():
        ids, labels = host_ids.to(dev, non_blocking=True), host_labels.to(dev, non_blocking=True)
        loss = engine(ids, labels)
        engine.backward(loss)
        engine.step()
        last["loss"] = loss.item()

    l0 = native.launch_count
    ms, clocks = _timed(torch, dist, step, args.steps, args.warmup)
    launches = native.launch_count - l0
    if rank == 0:
        from deepspeedsyclsupport_b200.sequence import fused as UF
        flops_tok = model.flops_per_token(S, with_recompute=False)
        extra = {"e2e": {"value": S / ms * 1e3, "unit": "tokens/s", "h2d_bytes_per_step": 2 * S_local * 8, "d2h_bytes_per_step": 4},
                 "final_loss": last.get("loss"), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
                 "model_flops_per_token": flops_tok, "tflops_per_gpu": round(S / ms * 1e3 * flops_tok / world / 1e12, 1),
                 "ulysses_transport": "peer heads-exchange kernels" if UF._Exchange._by_group else "nccl all-to-all",
                 "reference_arm": "not run (HF Llama + reference DistributedAttention at 128K needs flash-attn inside the reference stack)"}
        config = {"model": "llama-2-13b" + (f"-{args.layers}layers-INVALID-debug" if args.layers else ""), "global_batch": 1, "seq_len": S,
                  "parallelism": f"ulysses-sp{world}-zero3", "heads_per_rank": cfg.num_attention_heads // world, "optimizer": "AdamW",
                  "grad_clip": 1.0, "activation_checkpointing": getattr(model.model, "_policy_cache", (None, "auto"))[1],
                  "l2_policy": "working set far exceeds the 126 MB L2; no explicit flush"}
        print(json.dumps(_line("tokens/sec/box Llama-2-13B Ulysses sequence-parallel ZeRO-3 training step at long context", S / ms * 1e3, "tokens/s",
                               world, args, ms, config, clocks, launches, extra)), flush=True)


WORKLOADS = {"fastgen70b": run_fastgen70b, "mixtral": run_mixtral, "ulysses13b": run_ulysses13b}
