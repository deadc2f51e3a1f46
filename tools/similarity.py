"""Report difflib similarity of every package file against the same-named reference file (self-check for accidental
near-copies).  Usage: python tools/similarity.py [threshold]"""
import difflib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def norm(path):
    out = []
    for l in open(path, errors="ignore"):
        l = l.strip()
        if l and not l.startswith("#"):
            out.append(l)
    return out


def main():
    thr = float(sys.argv[1]) if len(sys.argv) > 1 else 0.5
    ref_by_name = {}
    for d, _, fs in os.walk(REF):
        if "/.git" in d:
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                ref_by_name.setdefault(f, []).append(os.path.join(d, f))
    rows = []
    for base in ("deepspeedsyclsupport_b200", "csrc", "bin", "tests"):
        for d, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f in ref_by_name and f != "__init__.py":
                    mine = norm(os.path.join(d, f))
                    if len(mine) < 8:
                        continue
                    for r in ref_by_name[f]:
                        theirs = norm(r)
                        ratio = difflib.SequenceMatcher(None, mine, theirs, autojunk=False).ratio()
                        if ratio >= thr:
                            rows.append((ratio, os.path.relpath(os.path.join(d, f), ROOT), os.path.relpath(r, REF)))
    for ratio, a, b in sorted(rows, reverse=True):
        print(f"{ratio:.2f}  {a}  ~  {b}")
    print(f"{len(rows)} file pairs at or above {thr}")


if __name__ == "__main__":
    main()
