"""Share of a file's code lines that sit inside runs of >= 5 consecutive lines identical to the same-path (or same-named)
reference file, and the longest such run.  Self-check against accidental near-copies.  Usage: python tools/copyrun.py [min_share]"""
import difflib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
PKG = "deepspeedsyclsupport_b200"


def norm(path):
    out = []
    for l in open(path, errors="ignore"):
        l = l.strip()
        if l and not l.startswith("#"):
            out.append(l.replace("'", '"'))
    return out


def runs(mine, theirs, k=5):
    sm = difflib.SequenceMatcher(None, mine, theirs, autojunk=False)
    blocks = [b.size for b in sm.get_matching_blocks() if b.size >= k]
    return sum(blocks), max(blocks, default=0)


def main():
    thr = float(sys.argv[1]) if len(sys.argv) > 1 else 0.15
    by_name = {}
    for d, _, fs in os.walk(os.path.join(REF, "deepspeed")):
        for f in fs:
            if f.endswith(".py"):
                by_name.setdefault(f, []).append(os.path.join(d, f))
    rows = []
    for d, _, fs in os.walk(os.path.join(ROOT, PKG)):
        for f in fs:
            if not f.endswith(".py") or f == "__init__.py":
                continue
            path = os.path.join(d, f)
            mine = norm(path)
            if len(mine) < 12:
                continue
            rel = os.path.relpath(path, os.path.join(ROOT, PKG))
            cands = [os.path.join(REF, "deepspeed", rel)] if os.path.exists(os.path.join(REF, "deepspeed", rel)) else by_name.get(f, [])
            best = (0, 0, None)
            for c in cands:
                tot, longest = runs(mine, norm(c))
                if tot > best[0]:
                    best = (tot, longest, c)
            if best[0] / len(mine) >= thr or best[1] >= 10:
                rows.append((best[0] / len(mine), best[1], len(mine), rel))
    for share, longest, n, rel in sorted(rows, reverse=True):
        print(f"{share:.2f}  longest {longest:3d}  lines {n:4d}  {rel}")
    print(f"{len(rows)} files")


if __name__ == "__main__":
    main()
