"""Verify the ctypes signature table in ops/native.py against the extern "C" declarations in csrc/."""
import glob
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from deepspeedsyclsupport_b200.ops import native
    # importing op modules registers their signatures
    import importlib
    for m in ("gemm", "fused_ops", "flat_optim", "quantizer", "attention", "flash_attention", "ragged_ops", "moe_ops", "misc_ops", "symm"):
        try:
            importlib.import_module("deepspeedsyclsupport_b200.ops." + m)
        except ImportError:
            pass
    for m in ("parallel.symm", "parallel.fused", "inference.v2.kernels", "inference.v2.kernels.cutlass_ops", "runtime.comm.compressed", "ops.random_ltd.dropping_utils", "ops.spatial", "ops.deepspeed4science.evoformer_attn", "moe.peer_dispatch", "sequence.fused"):
        try:
            importlib.import_module("deepspeedsyclsupport_b200." + m)
        except ImportError:
            pass
    src = "".join(open(f).read() for f in glob.glob(os.path.join(ROOT, "csrc", "**", "*.cu"), recursive=True))
    bad = 0
    exported = set(re.findall(r"DSB_EXPORT\s+[\w ]+?[\s\*]+(dsb200_\w+)\s*\(", src))
    for name, sig in native._SIGS.items():
        m = re.search(r"DSB_EXPORT\s+[\w ]+?\s+" + name + r"\s*\(([^)]*)\)", src, re.S)
        if not m:
            print("MISSING in csrc:", name)
            bad += 1
            continue
        params = [p.strip() for p in m.group(1).split(",") if p.strip()]
        codes = sig.replace(" ", "")
        ok = len(params) == len(codes)
        for p_, c in zip(params, codes):
            isptr = "*" in p_ or "cudaStream_t" in p_
            t = "p" if isptr else ("l" if "long long" in p_ else ("f" if "float" in p_ else "i"))
            if t != c:
                ok = False
                print("  mismatch", name, repr(p_), "declared", c)
        if not ok:
            print("BAD", name, len(params), len(codes))
            bad += 1
    for name in sorted(exported - set(native._SIGS) - {"dsb200_last_error", "dsb200_version"}):
        print("no python signature for", name)
    print("signature check:", "FAILED" if bad else "ok", f"({len(native._SIGS)} entry points)")
    return bad


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
