"""Summarise an .ncu-rep (raw page) and the SASS of libdsb200.so into profiles/ (tracked evidence)."""
import csv
import io
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.avg.per_second"]


def ncu_summary(rep, out_md, title):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    lines = [f"# {title}", "", f"source: `{os.path.basename(rep)}` (ncu --set full --clock-control none --import-source on), B200", ""]
    seen = Counter()
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        short = re.sub(r"\(.*", "", name)
        seen[short] += 1
        if seen[short] > 1:
            continue
        lines.append(f"## {short}")
        lines.append("")
        lines.append("| metric | value | unit |")
        lines.append("|---|---|---|")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                lines.append(f"| {w} | {r[i]} | {units[i]} |")
        lines.append("")
    open(out_md, "w").write("\n".join(lines) + "\n")


def sass_summary(out_md):
    so = os.path.join(ROOT, "deepspeedsyclsupport_b200", "lib", "libdsb200.so")
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    cur = None
    per = {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            for key in ("UTCHMMA", "UTCQMMA", "UTCCP", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "LDTM", "STTM", "UTCATOMSWS", "HMMA", "SYNCS",
                        "LDG", "STG", "LDGSTS", "MUFU", "ATOM", "RED"):
                if op.startswith(key):
                    per[cur][key] += 1
            if "MC" in op.split(".")[0] or ".MC" in op or "LDGMC" in op:
                per[cur]["MULTIMEM"] += 1
            if op.startswith("UTCHMMA") and ".2CTA" in op:
                per[cur]["2CTA"] += 1
    lines = ["# SASS evidence (cuobjdump -sass libdsb200.so, sm_100a)", "",
             "Counts of Blackwell-specific mnemonics per kernel: `UTCHMMA` = tcgen05.mma (bf16; `.2CTA` = cta_group::2 pair MMA), `UTMALDG/UTMASTG/UTMAREDG` = TMA "
             "load/store/reduce, `LDTM/STTM` = tcgen05.ld/st, `SYNCS` = mbarrier ops, `HMMA` = legacy mma.sync (must be 0).", "",
             "`UTCQMMA` = tcgen05.mma block-scaled (mxf8f6f4.block_scale), `UTCCP` = tcgen05.cp (smem -> TMEM), `UBLKCP` = cp.async.bulk, "
             "`MULTIMEM` = multimem.ld_reduce / multimem.st (NVLS: LDGMC / STG...MC forms).", "",
             "| kernel | UTCHMMA | of which .2CTA | UTCQMMA | UTCCP | UTMALDG | UTMASTG | UTMAREDG | UBLKCP | LDTM | STTM | SYNCS | MULTIMEM | HMMA | LDG | STG |",
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for k, c in sorted(per.items()):
        if sum(c.values()) == 0:
            continue
        dem = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
        dem = re.sub(r"\(.*", "", dem)[:90]
        lines.append(f"| `{dem}` | {c['UTCHMMA']} | {c['2CTA']} | {c['UTCQMMA']} | {c['UTCCP']} | {c['UTMALDG']} | {c['UTMASTG']} | {c['UTMAREDG']} | {c['UBLKCP']} | "
                     f"{c['LDTM']} | {c['STTM']} | {c['SYNCS']} | {c['MULTIMEM']} | {c['HMMA']} | {c['LDG']} | {c['STG']} |")
    open(out_md, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    if len(sys.argv) > 1:
        ncu_summary(sys.argv[1], os.path.join(ROOT, "profiles", sys.argv[2]), sys.argv[3] if len(sys.argv) > 3 else "ncu summary")
    sass_summary(os.path.join(ROOT, "profiles", "sass_evidence.md"))
