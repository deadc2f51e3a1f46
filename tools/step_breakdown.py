"""Per-kernel time breakdown of one flagship training step (Llama ZeRO-3 bf16) with the torch profiler (CUPTI).
Diagnostic only -- numbers under a profiler are never benchmark values.  Usage:
    python tools/step_breakdown.py [--layers N] [--micro-batch B] [--seq-len S] [--out profiles/step_breakdown.txt]"""
import argparse
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--model", default="7b")
    p.add_argument("--layers", type=int, default=None)
    p.add_argument("--micro-batch", type=int, default=2)
    p.add_argument("--seq-len", type=int, default=4096)
    p.add_argument("--out", default=None)
    a = p.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    for k, v in (("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_SIZE", "1")):
        os.environ.setdefault(k, v)
    import torch
    import bench
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    deepspeed.init_distributed()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    over = {"activation_checkpointing": True}
    if a.layers:
        over["num_hidden_layers"] = a.layers
    cfg = llama_config(a.model, **over)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg)
    torch.set_default_dtype(torch.float32)
    args = argparse.Namespace(micro_batch=a.micro_batch, seq_len=a.seq_len, zero_stage=3, no_act_ckpt=False, gpus=1)
    engine, _, _, _ = deepspeed.initialize(model=model, config=bench.ds_config(args, world))
    ids = torch.randint(0, cfg.vocab_size, (a.micro_batch, a.seq_len), device=dev)

    def step():
        loss = engine(ids, ids)
        engine.backward(loss)
        engine.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    step()
    t1.record()
    torch.cuda.synchronize()
    plain_ms = t0.elapsed_time(t1)
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
        torch.cuda.synchronize()
    if rank != 0:
        return
    agg = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            name = ev.name
            agg[name][0] += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
            agg[name][1] += 1
    total = sum(v[0] for v in agg.values())
    lines = [f"model={a.model} layers={cfg.num_hidden_layers} mb={a.micro_batch} seq={a.seq_len}: sum of GPU kernel time {total / 1e3:.1f} ms (one step, under profiler); step without profiler {plain_ms:.1f} ms; world {world}",
             f"{'kernel':<90} {'calls':>6} {'ms':>9} {'%':>6}"]
    for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
        lines.append(f"{name[:90]:<90} {n:>6} {us / 1e3:>9.2f} {100 * us / total:>6.1f}")
    text = "\n".join(lines)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
