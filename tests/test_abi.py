"""C ABI hygiene: the public header declares every exported entry point, compiles as C, and the ctypes signature table
matches the extern "C" definitions."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_is_complete_and_valid_c():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_header.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run(["gcc", "-fsyntax-only", "-x", "c", os.path.join(ROOT, "include", "dsb200.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_ctypes_signatures_match_sources():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_sigs.py")], capture_output=True, text=True)
    assert r.returncode == 0 and "no python signature" not in r.stdout, r.stdout + r.stderr
