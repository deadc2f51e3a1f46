"""Do the ZeRO-3 gather copies run UNDER compute?  One profiled training step (torch profiler / CUPTI); for every peer copy or
collective kernel prints when it started, how long it ran, and how much of that time some compute kernel of another stream was
running.  Diagnostic only (profiler on).  torchrun ... tools/overlap_probe.py [--layers N]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    for k, v in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29511"), ("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1")):
        os.environ.setdefault(k, v)
    import torch
    import bench
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    deepspeed.init_distributed()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    cfg = llama_config("7b", activation_checkpointing=True, num_hidden_layers=a.layers)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg)
    torch.set_default_dtype(torch.float32)
    args = argparse.Namespace(micro_batch=2, seq_len=4096, zero_stage=3, no_act_ckpt=False, gpus=world)
    engine, _, _, _ = deepspeed.initialize(model=model, config=bench.ds_config(args, world))
    ids = torch.randint(0, cfg.vocab_size, (2, 4096), device=dev)

    def step():
        loss = engine(ids, ids)
        engine.backward(loss)
        engine.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    if rank != 0:
        return
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    comm_tags = ("Memcpy PtoP", "symm::", "nccl")
    comm = [e for e in evs if any(t in e.name for t in comm_tags)]
    comp = sorted((e for e in evs if not any(t in e.name for t in comm_tags) and "Memcpy" not in e.name and "Memset" not in e.name),
                  key=lambda e: e.time_range.start)
    t0 = min(e.time_range.start for e in evs)
    lines = [f"world {world}, {a.layers} layers; times in ms from the first GPU activity of the step", f"{'comm op':<60} {'start':>9} {'dur':>8} {'under compute':>14}"]
    for e in sorted(comm, key=lambda e: e.time_range.start):
        s, f = e.time_range.start, e.time_range.end
        covered = sum(max(0, min(f, c.time_range.end) - max(s, c.time_range.start)) for c in comp if c.time_range.end > s and c.time_range.start < f)
        lines.append(f"{e.name[:60]:<60} {(s - t0) / 1e3:>9.3f} {(f - s) / 1e3:>8.3f} {min(1.0, covered / max(1, f - s)):>13.0%}")
    text = "\n".join(lines)
    print(text)
    if a.out:
        with open(a.out, "w") as fd:
            fd.write(text + "\n")


if __name__ == "__main__":
    main()
