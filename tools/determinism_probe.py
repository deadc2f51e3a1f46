"""Is one engine backward reproducible?  Runs forward+backward N times from the same weights (no optimizer step in between) and
prints the relative difference of the reduced gradient shard between repetitions, per unit.  Diagnostic.
  python tools/determinism_probe.py            (world 1)
  torchrun --nproc-per-node 2 ... tools/determinism_probe.py [--no-p2p] [--no-overlap]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-p2p", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--instances", type=int, default=2, help="engines built one after the other in this process (same seed, same data)")
    a = ap.parse_args()
    for k, v in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29512"), ("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1")):
        os.environ.setdefault(k, v)
    if a.no_p2p:
        os.environ["DSB200_DISABLE_P2P"] = "1"
    import torch
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    deepspeed.init_distributed()
    rank = dist.get_rank()
    first = []
    for inst in range(a.instances):
        torch.manual_seed(0)
        cfg = llama_config("tiny", hidden_size=256, intermediate_size=768, num_attention_heads=2, num_key_value_heads=2,
                           num_hidden_layers=3, vocab_size=1024, activation_checkpointing=False)
        model = LlamaForCausalLM(cfg).cuda().bfloat16()
        engine, _, _, _ = deepspeed.initialize(model=model, config={
            "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "gradient_clipping": 1.0,
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
            "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "overlap_comm": not a.no_overlap}})
        opt = engine.optimizer
        weights = torch.cat([opt.get_full_fp32_param(p).float().flatten() for p in model.parameters()])
        g = torch.Generator().manual_seed(5 + rank)
        ids = torch.randint(0, 1024, (2, a.seq), generator=g).cuda()
        labels = torch.randint(0, 1024, (2, a.seq), generator=g).cuda()
        snaps, losses = [], []
        for r in range(a.reps):
            loss = engine(ids, labels)
            engine.backward(loss)
            if hasattr(opt, "rs_stream") and opt.rs_stream is not None:
                torch.cuda.current_stream().wait_stream(opt.rs_stream)
            torch.cuda.synchronize()
            snaps.append(opt.grad_shard.float().clone())
            losses.append(loss.item())
            opt.zero_grad()
            if opt.p2p is not None:  # what step() does to the slot ring, without updating anything
                opt._p2p_next_slot = 0
                opt._p2p_reduce_count = 0
                opt._p2p_barrier_events = {}
                opt._p2p_sumsq.zero_()
            torch.cuda.synchronize()
            dist.barrier()
        first.append((snaps[0], weights, losses[0], ids.clone()))
        if rank == 0:
            mode = "nccl" if opt.p2p is None else f"p2p({opt._p2p_rs_mode})"
            print(f"instance {inst}: world {dist.get_world_size()} path {mode} overlap {not a.no_overlap}: losses {losses}")
            for r in range(1, a.reps):
                rel = ((snaps[r] - snaps[0]).norm() / snaps[0].norm()).item()
                per = []
                for u in opt.units:
                    s, e = u.shard_off, u.shard_off + u.shard_numel
                    per.append(f"{u.name}={((snaps[r][s:e] - snaps[0][s:e]).norm() / (snaps[0][s:e].norm() + 1e-20)).item():.1e}")
                print(f"  rep {r} vs 0: {rel:.3e}   " + " ".join(per))
    if rank == 0 and len(first) > 1:
        for i in range(1, len(first)):
            print(f"instance {i} vs 0: grads rel diff {((first[i][0] - first[0][0]).norm() / first[0][0].norm()).item():.3e}  "
                  f"weights max diff {(first[i][1] - first[0][1]).abs().max().item():.3e}  loss {first[i][2]!r} vs {first[0][2]!r}  "
                  f"same ids {bool((first[i][3] == first[0][3]).all())}")


if __name__ == "__main__":
    main()
