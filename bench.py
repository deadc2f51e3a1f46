#!/usr/bin/env python
"""Headline benchmark: tokens/s/box, Llama-2-7B, ZeRO-3, bf16, AdamW, activation checkpointing, synthetic data.

  python bench.py --gpus N --steps K --warmup W            # this framework (own kernels + engine)
  python bench.py --impl reference --gpus N ...            # UNMODIFIED reference DeepSpeed 0.12.7 from baseline/_ref
                                                           # driving HF LlamaForCausalLM through deepspeed.initialize()

For N > 1 launch with torchrun (one rank per GPU).  Rank 0 prints ONE JSON line.  Both arms run the same model shape,
micro-batch, sequence length, optimizer, precision, ZeRO stage and activation-checkpointing setting.
`value` is device-timed (CUDA events, max over ranks) with inputs staged on the device; `e2e.value` is the same loop
through the public API with a pinned-host -> device copy of every step's tokens and a device -> host read of the loss.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="own", choices=["own", "reference"])
    p.add_argument("--model", default="7b")
    p.add_argument("--layers", type=int, default=None, help="debug only: override layer count (number becomes INVALID)")
    p.add_argument("--micro-batch", type=int, default=2)
    p.add_argument("--seq-len", type=int, default=4096)
    p.add_argument("--no-act-ckpt", action="store_true")
    p.add_argument("--zero-stage", type=int, default=3)
    p.add_argument("--zeropp", action="store_true", help="ZeRO++: zero_quantized_weights + zero_quantized_gradients (config #2, second half)")
    p.add_argument("--workload", default="llama7b", choices=["llama7b", "fastgen70b", "mixtral", "ulysses13b"],
                   help="llama7b = the headline (config #2); the others are the end-to-end runs of configs #5 / #4 / #3 (bench_workloads.py)")
    p.add_argument("--quant", default="", help="fastgen70b: '' (bf16) | wf8af8_mx | wf8af8_block | wf8af8")
    p.add_argument("--sequences", type=int, default=64, help="fastgen70b: concurrent sequences")
    p.add_argument("--prompt-min", type=int, default=256)
    p.add_argument("--prompt-max", type=int, default=1024)
    p.add_argument("--token-budget", type=int, default=2048, help="fastgen70b: tokens per forward (ragged batch size)")
    p.add_argument("--gen-tokens", type=int, default=16, help="fastgen70b: decode steps per timed step")
    p.add_argument("--local_rank", type=int, default=-1)
    p.add_argument("--ref-child", action="store_true", help="internal: run one reference configuration in this process")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi clock / throttle sampling during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index=0):
        self.proc = None
        self.lines = []
        self.idx = device_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ds_config(args, world):
    cfg = _ds_config(args, world)
    if getattr(args, "zeropp", False):
        cfg["zero_optimization"].update({"zero_quantized_weights": True, "zero_quantized_gradients": True})
    return cfg


def _ds_config(args, world):
    return {
        "train_micro_batch_size_per_gpu": args.micro_batch,
        "gradient_accumulation_steps": 1,
        "steps_per_print": 1000000,
        "bf16": {"enabled": True},
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
        "gradient_clipping": 1.0,
        "zero_optimization": {"stage": args.zero_stage, "overlap_comm": True, "contiguous_gradients": True,
                              "stage3_prefetch_bucket_size": 5e8, "stage3_max_live_parameters": 1e10,  # 180 GB GPUs: gathered bf16 weights stay live fwd -> bwd
                              
                              "stage3_param_persistence_threshold": 1e5, "reduce_bucket_size": 5e8},
    }


def result_line(args, impl, world, tokens_per_step, ms_own, e2e_tps, h2d, d2h, launches, clocks, extra=None, act_ckpt=False):
    value = tokens_per_step / (ms_own / 1e3)
    out = {
        "metric": "tokens/sec/box Llama-2-7B ZeRO-3 bf16 (training step: fwd+bwd+AdamW)",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_own, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic tokens (uniform random ids), random-init weights", "impl": impl,
        "config": {"model": f"llama-2-{args.model}" + (f"-{args.layers}layers-INVALID-debug" if args.layers else ""),
                   "global_batch": args.micro_batch * world, "seq_len": args.seq_len, "micro_batch_per_gpu": args.micro_batch,
                   "parallelism": f"zero{args.zero_stage}{'++(qwZ+qgZ)' if args.zeropp else ''}-dp{world}", "optimizer": "AdamW", "grad_clip": 1.0,
                   # resolved value, same vocabulary in both arms: False = nothing is recomputed, else the recompute policy
                   "activation_checkpointing": act_ckpt,
                   "l2_policy": "working set (weights+activations, >13 GB) far exceeds the 126 MB L2; no explicit flush"},
        "e2e": {"value": e2e_tps, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "clocks": clocks,
    }
    if extra:
        out.update(extra)
    return out


def timed_loops(args, step_fn, step_fn_e2e, world, dist_mod, torch, count_fn):
    """W warm-up steps, K device-timed steps, then K end-to-end steps. Returns (ms_per_step, e2e_tokens_per_s, launches, clocks)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    for _ in range(args.warmup):
        step_fn()
    torch.cuda.synchronize()
    sampler = ClockSampler(torch.cuda.current_device())
    if int(os.environ.get("RANK", "0")) == 0:
        sampler.start()
    dist_mod.barrier()
    torch.cuda.synchronize()
    l0 = count_fn()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(args.steps):
        step_fn()
    end.record()
    torch.cuda.synchronize()
    dist_mod.barrier()
    torch.cuda.synchronize()
    launches = count_fn() - l0
    ms = torch.tensor([start.elapsed_time(end)], device=dev)
    dist_mod.all_reduce(ms, op=dist_mod.ReduceOp.MAX)
    ms_per_step = ms.item() / args.steps
    # ---- end to end: pinned host tokens -> device every step, loss read back every step
    dist_mod.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_fn_e2e()
    torch.cuda.synchronize()
    dist_mod.barrier()
    dt = torch.tensor([time.perf_counter() - t0], device=dev)
    dist_mod.all_reduce(dt, op=dist_mod.ReduceOp.MAX)
    clocks = sampler.stop() if int(os.environ.get("RANK", "0")) == 0 else None
    tokens_per_step = world * args.micro_batch * args.seq_len
    return ms_per_step, tokens_per_step * args.steps / dt.item(), launches, clocks


def run_own(args):
    sys.path.insert(0, ROOT)
    import torch
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    from deepspeedsyclsupport_b200.ops import native
    deepspeed.init_distributed()
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    over = {"activation_checkpointing": not args.no_act_ckpt}
    if args.layers:
        over["num_hidden_layers"] = args.layers
    cfg = llama_config(args.model, **over)
    torch.manual_seed(1234)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg)
    torch.set_default_dtype(prev)
    engine, _, _, _ = deepspeed.initialize(model=model, config=ds_config(args, world))
    B, S, V = args.micro_batch, args.seq_len, cfg.vocab_size
    g = torch.Generator().manual_seed(1000 + rank)
    host_ids = torch.randint(0, V, (B, S), generator=g).pin_memory()
    host_labels = torch.randint(0, V, (B, S), generator=g).pin_memory()
    dev_ids, dev_labels = host_ids.to(dev), host_labels.to(dev)
    last = {}

    def step():
        loss = engine(dev_ids, dev_labels)
        engine.backward(loss)
        engine.step()
        last["loss"] = loss

    def step_e2e():
        ids = host_ids.to(dev, non_blocking=True)
        labels = host_labels.to(dev, non_blocking=True)
        loss = engine(ids, labels)
        engine.backward(loss)
        engine.step()
        last["loss_host"] = loss.item()

    ms, e2e, launches, clocks = timed_loops(args, step, step_e2e, world, dist, torch, lambda: native.launch_count)
    # exposed communication: separate pass with the optimizer's comm probe on (not inside the timed region)
    exposed_ms, exposed_sites, nvlink_rx = 0.0, {}, 0
    opt = engine.optimizer
    if world > 1 and hasattr(opt, "comm_probe_start"):
        k = max(2, min(args.steps, 4))
        torch.cuda.synchronize()
        dist.barrier()
        opt.comm_probe_start()
        rx0 = getattr(opt, "nvlink_rx_bytes", 0)
        for _ in range(k):
            step()
        nvlink_rx = (getattr(opt, "nvlink_rx_bytes", 0) - rx0) // k
        total, sites = opt.comm_probe_stop(detail=True)
        t = torch.tensor([total / k], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        exposed_ms = round(t.item(), 3)
        exposed_sites = {s_: round(v / k, 3) for s_, v in sorted(sites.items())}  # rank 0's split, for orientation
    if rank == 0:
        h2d = host_ids.numel() * host_ids.element_size() + host_labels.numel() * host_labels.element_size()
        extra = {"final_loss": last.get("loss_host"), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
                 "model_flops_per_token": model.flops_per_token(S, with_recompute=False)}  # 6*P + attention, no recompute credit
        tps = world * B * S / (ms / 1e3)
        extra["tflops_per_gpu"] = round(tps / world * extra["model_flops_per_token"] / 1e12, 1)
        resolved = getattr(model.model, "_policy_cache", (None, None))[1]
        act = False if (args.no_act_ckpt or resolved in (None, "none")) else resolved
        extra["exposed_comm_ms"] = exposed_ms
        extra["exposed_comm_sites_rank0"] = exposed_sites
        extra["nvlink_rx_bytes_per_step_per_gpu"] = int(nvlink_rx)  # payload bytes of the ZeRO collectives, counted by the optimizer
        if os.environ.get("DSB200_PROBE_TRACE") == "1":
            extra["exposed_comm_trace_rank0"] = [ms for site, ms in getattr(opt, "comm_probe_trace", []) if site == "gather_wait"]
        extra["exposed_comm_how"] = ("CUDA-event pairs around every point where the compute stream waits for a gather / reduce stream or "
                                     "runs a collective itself (grad-norm all-reduce, post-step barrier + persistent gathers); "
                                     "per step, max over ranks, measured in a separate pass after the timed one")
        extra["zero_collectives"] = getattr(engine.optimizer, "_p2p_rs_mode", None) and {
            "all_gather": engine.optimizer._p2p_ag_mode, "reduce_scatter": engine.optimizer._p2p_rs_mode,
            "symm_backend": engine.optimizer.p2p.backend} or "nccl"
        line = result_line(args, "own", world, world * B * S, ms, e2e, h2d, 4, launches, clocks, extra, act_ckpt=act)
        print(json.dumps(line), flush=True)


def run_reference(args):
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import ref_compat
        ref_dir = ref_compat.install_shims()
        if not os.path.isdir(os.path.join(ref_dir, "deepspeed")):
            raise RuntimeError("baseline/_ref/deepspeed not installed (pip install --no-deps --target baseline/_ref /root/reference)")
        import torch
        import deepspeed
        from transformers import LlamaConfig, LlamaForCausalLM
        # the reference's init_distributed() does not select a device (its engine does, later): do it here, the way its
        # launcher-driven examples do with args.local_rank, so the model and the inputs of rank r live on GPU r
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        deepspeed.init_distributed()
        import torch.distributed as tdist
        world, rank = tdist.get_world_size(), tdist.get_rank()
        dev = torch.device("cuda", torch.cuda.current_device())
        shapes = {"7b": dict(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32),
                  "13b": dict(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=40),
                  "tiny": dict(hidden_size=128, intermediate_size=384, num_hidden_layers=2, num_attention_heads=2, num_key_value_heads=2, vocab_size=512)}[args.model]
        if args.layers:
            shapes["num_hidden_layers"] = args.layers
        hf_cfg = LlamaConfig(vocab_size=shapes.pop("vocab_size", 32000), max_position_embeddings=max(4096, args.seq_len),
                             rms_norm_eps=1e-5, tie_word_embeddings=False, attn_implementation="sdpa", use_cache=False, **shapes)
        torch.manual_seed(1234)
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        with torch.device(dev):
            model = LlamaForCausalLM(hf_cfg)
        torch.set_default_dtype(prev)
        # N == 1 runs both variants (run_reference_best_of); at N > 1 the sharded model leaves even more room, so the arm
        # runs the variant that won at N == 1 -- no recomputation, the same setting the own arm resolves to
        use_ckpt = not args.no_act_ckpt and world == 1
        if use_ckpt:
            model.gradient_checkpointing_enable()
        model.train()
        cfg = ds_config(args, world)
        notes = []
        try:
            engine, _, _, _ = deepspeed.initialize(model=model, config=cfg, model_parameters=[p for p in model.parameters() if p.requires_grad])
        except Exception as e:  # FusedAdam JIT build can fail offline: fall back to the reference's own torch_adam switch
            notes.append(f"FusedAdam unavailable ({type(e).__name__}: {str(e)[:120]}); reference run with torch_adam=true")
            cfg["optimizer"]["params"]["torch_adam"] = True
            engine, _, _, _ = deepspeed.initialize(model=model, config=cfg, model_parameters=[p for p in model.parameters() if p.requires_grad])
        B, S, V = args.micro_batch, args.seq_len, hf_cfg.vocab_size
        g = torch.Generator().manual_seed(1000 + rank)
        host_ids = torch.randint(0, V, (B, S), generator=g).pin_memory()
        host_labels = torch.randint(0, V, (B, S), generator=g).pin_memory()
        dev_ids, dev_labels = host_ids.to(dev), host_labels.to(dev)
        last = {}
        # Watchdog: if a rank makes no progress for DSB_REF_WATCHDOG seconds (e.g. a peer died and left this one inside a
        # collective), dump every thread's stack and report the arm as unavailable instead of hanging the caller.
        import faulthandler
        import threading
        limit = float(os.environ.get("DSB_REF_WATCHDOG", "180"))
        progress = {"t": time.time(), "done": False}

        def _watch():
            while not progress["done"]:
                time.sleep(5)
                if time.time() - progress["t"] > limit:
                    faulthandler.dump_traceback(file=sys.stderr, all_threads=True)
                    if rank == 0:
                        print(json.dumps({"impl": "reference", "n_gpus": world,
                                          "unavailable": f"reference training step made no progress for {int(limit)} s at world size {world} (see the stack dump on stderr)"}), flush=True)
                    os._exit(0)

        threading.Thread(target=_watch, daemon=True).start()

        def step():
            progress["t"] = time.time()
            loss = engine(input_ids=dev_ids, labels=dev_labels).loss
            engine.backward(loss)
            engine.step()

        def step_e2e():
            progress["t"] = time.time()
            ids = host_ids.to(dev, non_blocking=True)
            labels = host_labels.to(dev, non_blocking=True)
            loss = engine(input_ids=ids, labels=labels).loss
            engine.backward(loss)
            engine.step()
            last["loss_host"] = loss.item()

        class _D:  # torch.distributed adapter with the names timed_loops uses
            barrier = staticmethod(tdist.barrier)
            all_reduce = staticmethod(tdist.all_reduce)
            ReduceOp = tdist.ReduceOp

        ms, e2e, _, clocks = timed_loops(args, step, step_e2e, world, _D, torch, lambda: 0)
        progress["done"] = True
        if rank == 0:
            h2d = host_ids.numel() * 8 * 2
            extra = {"final_loss": last.get("loss_host"), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
                     "reference": "DeepSpeed 0.12.7 (baseline/_ref, unmodified) + transformers LlamaForCausalLM sdpa", "notes": notes}
            line = result_line(args, "reference", world, world * B * S, ms, e2e, h2d, 4, 0, clocks, extra,
                               act_ckpt="full" if use_ckpt else False)
            print(json.dumps(line), flush=True)
    except Exception as e:
        import traceback
        traceback.print_exc(file=sys.stderr)  # every rank: a failure on rank > 0 must not look like a hang of rank 0
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}), flush=True)
        sys.exit(0)


def main():
    args = parse()
    if os.environ.get("DSB200_HANG_DUMP"):  # diagnosis: dump every thread's stack if the process is still running after N seconds
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["DSB200_HANG_DUMP"]), repeat=False, file=sys.stderr)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    # N == 1 is launched without torchrun: behave like a one-rank job (the reference would otherwise probe for MPI)
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("LOCAL_SIZE", os.environ["WORLD_SIZE"])
    if args.workload != "llama7b":
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": f"no reference arm for --workload {args.workload} (see the line's reference_arm note)"}))
            return
        import bench_workloads
        return bench_workloads.WORKLOADS[args.workload](args)
    if args.impl == "reference":
        if int(os.environ["WORLD_SIZE"]) == 1 and not args.ref_child and not args.no_act_ckpt:
            run_reference_best_of(args)
        else:
            run_reference(args)
    else:
        run_own(args)


def run_reference_best_of(args):
    """The own arm keeps activations instead of recomputing them when memory allows; to compare like with like the
    reference gets the same freedom: it is run (unmodified, each time in a fresh process) once without and once with
    HF gradient checkpointing, and the faster configuration that works is reported (both results are listed in `tried`)."""
    base = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--ref-child", "--gpus", "1", "--steps", str(args.steps),
            "--warmup", str(args.warmup), "--model", args.model, "--micro-batch", str(args.micro_batch), "--seq-len", str(args.seq_len),
            "--zero-stage", str(args.zero_stage)] + (["--layers", str(args.layers)] if args.layers else []) + (["--zeropp"] if args.zeropp else [])
    tried, best = [], None
    for extra in (["--no-act-ckpt"], []):
        line = None
        try:
            r = subprocess.run(base + extra, capture_output=True, text=True, timeout=900)
            for l in r.stdout.splitlines():
                if l.startswith("{"):
                    line = json.loads(l)
            sys.stderr.write(r.stderr[-4000:])
        except subprocess.TimeoutExpired:
            line = {"impl": "reference", "unavailable": "timeout"}
        label = "no gradient checkpointing" if extra else "gradient checkpointing"
        if line is None:
            line = {"impl": "reference", "unavailable": "no result line (crash / out of memory)"}
        tried.append({"variant": label, "ms_per_step": line.get("ms_per_step"), "unavailable": line.get("unavailable")})
        if "value" in line and (best is None or line["value"] > best["value"]):
            best = line
    if best is None:
        print(json.dumps({"impl": "reference", "unavailable": "; ".join(f"{t['variant']}: {t['unavailable']}" for t in tried)}), flush=True)
        return
    best["tried"] = tried
    print(json.dumps(best), flush=True)


if __name__ == "__main__":
    main()
