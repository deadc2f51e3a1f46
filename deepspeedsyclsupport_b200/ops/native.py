"""ctypes bridge to the in-tree native libraries.

`lib()` returns libdsb200.so (CUDA, sm_100a); `host_lib()` returns libdsb200_host.so (CPU).
Ops fail loudly when a GPU is present but the device library is missing -- there is no silent
PyTorch fallback on a GPU box (the pure-PyTorch reference implementations in `ops/*` exist for
CPU-only plumbing tests and as numerical oracles).
"""
import ctypes
import os
import threading

import torch

_LIB_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib")
_lock = threading.Lock()
_dev = None
_host = None
launch_count = 0  # number of kernels of OUR library launched (bench.py reports it as gpu_launches)


class NativeError(RuntimeError):
    pass


def _load(name, build_fn):
    path = os.path.join(_LIB_DIR, name)
    if not os.path.exists(path):
        try:
            from .op_builder import builder
            getattr(builder, build_fn)()
        except Exception as e:  # pragma: no cover
            raise NativeError(f"{name} is not built and building it failed: {e}")
    if not os.path.exists(path):
        raise NativeError(f"{name} missing; run `python -m deepspeedsyclsupport_b200.ops.op_builder.builder`")
    return ctypes.CDLL(path)


def lib():
    global _dev
    if _dev is None:
        with _lock:
            if _dev is None:
                l = _load("libdsb200.so", "build_device")
                l.dsb200_last_error.restype = ctypes.c_char_p
                _dev = l
    return _dev


def host_lib():
    global _host
    if _host is None:
        with _lock:
            if _host is None:
                l = _load("libdsb200_host.so", "build_host")
                if hasattr(l, "dsbh_last_error"):
                    l.dsbh_last_error.restype = ctypes.c_char_p
                _host = l
    return _host


def available():
    """True when CUDA kernels can run in this process."""
    return torch.cuda.is_available()


def check(rc, n_launch=1):
    global launch_count
    if rc != 0:
        raise NativeError(lib().dsb200_last_error().decode())
    launch_count += n_launch


def stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


c_int = ctypes.c_int
c_float = ctypes.c_float
c_long = ctypes.c_int64
c_void_p = ctypes.c_void_p
