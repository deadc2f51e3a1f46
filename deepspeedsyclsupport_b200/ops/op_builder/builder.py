"""Ahead-of-time builder for the native libraries (sm_100a only).

The reference has one OpBuilder subclass per op that JIT-compiles through
torch.utils.cpp_extension for whatever device is visible (op_builder/builder.py:462-604,
op_builder/all_ops.py:17-32).  This build targets exactly one architecture, so the whole
device library is one `nvcc -shared` invocation per translation unit, produced in-tree
(`deepspeedsyclsupport_b200/lib/libdsb200.so`) and loaded through its C ABI with ctypes;
the host-only natives (AVX-512 CPU optimizers, async file I/O, ragged-batch bookkeeping)
live in `libdsb200_host.so`, which builds and runs without a GPU.
"""
import glob
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

_PKG = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REPO_ROOT = os.path.dirname(_PKG)
CSRC = os.path.join(REPO_ROOT, "csrc")
LIB_DIR = os.path.join(_PKG, "lib")
BUILD_DIR = os.path.join(REPO_ROOT, "build")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--use_fast_math", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
]
HOST_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-fvisibility=hidden", "-march=x86-64-v3"]


def nvcc_path():
    cand = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    return cand if os.path.exists(cand) else shutil.which("nvcc")


def _digest(paths, extra=""):
    h = hashlib.sha1(extra.encode())
    for p in sorted(paths):
        h.update(p.encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def device_sources():
    return sorted(glob.glob(os.path.join(CSRC, "**", "*.cu"), recursive=True))


def host_sources():
    return sorted(glob.glob(os.path.join(CSRC, "host", "*.cpp")))


def _headers():
    return sorted(glob.glob(os.path.join(CSRC, "**", "*.cuh"), recursive=True) +
                  glob.glob(os.path.join(CSRC, "**", "*.h"), recursive=True) +
                  glob.glob(os.path.join(REPO_ROOT, "include", "*.h")))


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build_device(force=False, verbose=False):
    """Compile every csrc/**/*.cu into lib/libdsb200.so (incremental per translation unit)."""
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found")
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(BUILD_DIR, exist_ok=True)
    out = os.path.join(LIB_DIR, "libdsb200.so")
    hdr_digest = _digest(_headers(), " ".join(NVCC_FLAGS))
    objs, jobs = [], []
    for src in device_sources():
        rel = os.path.relpath(src, CSRC).replace(os.sep, "__")
        obj = os.path.join(BUILD_DIR, rel + ".o")
        stamp = obj + ".sha1"
        want = _digest([src], hdr_digest)
        have = open(stamp).read() if os.path.exists(stamp) and os.path.exists(obj) else ""
        objs.append(obj)
        if force or have != want:
            jobs.append((src, obj, stamp, want))

    def compile_one(job):
        src, obj, stamp, want = job
        log = _run([nvcc] + NVCC_FLAGS + ["-I", os.path.join(REPO_ROOT, "include"), "-c", src, "-o", obj])
        with open(stamp, "w") as f:
            f.write(want)
        if verbose and log.strip():
            print(log)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(compile_one, jobs))
    if jobs or not os.path.exists(out):
        _run([nvcc, "-shared", "-o", out] + objs + ["-lcudart"])
    return out


def build_host(force=False):
    """Compile csrc/host/*.cpp into lib/libdsb200_host.so (no CUDA needed)."""
    os.makedirs(LIB_DIR, exist_ok=True)
    out = os.path.join(LIB_DIR, "libdsb200_host.so")
    srcs = host_sources()
    if not srcs:
        return None
    stamp = out + ".sha1"
    want = _digest(srcs + _headers(), " ".join(HOST_FLAGS))
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == want:
        return out
    libs = []
    if os.path.exists("/usr/include/libaio.h"):
        libs = ["-DDSB_HAVE_LIBAIO", "-laio"]
    _run(["g++"] + HOST_FLAGS + ["-shared", "-o", out] + srcs + ["-lpthread"] + libs)
    with open(stamp, "w") as f:
        f.write(want)
    return out


def build_all(force=False, verbose=False):
    return build_device(force, verbose), build_host(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose=True))
