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


# C-ABI signatures (include/dsb200.h): p = pointer, i = int, l = long long, f = float.  Declaring argtypes makes ctypes
# convert plain Python numbers correctly (a bare Python float would otherwise be passed as a double).
_T = {"p": ctypes.c_void_p, "i": ctypes.c_int, "l": ctypes.c_longlong, "f": ctypes.c_float}
_SIGS = {
    "dsb200_gemm_bf16": "pppiiiiiiiiiip",
    "dsb200_moe_gemm_bf16": "pppp iiii iii p",
    "dsb200_gemm_fp8": "ppppp iii iii i p",
    "dsb200_flash_fwd_trace": "p",
    "dsb200_fused_adam_flat": "ppppipil fffff iii ppp",
    "dsb200_fused_lion_flat": "pp pipil ffff ppp",
    "dsb200_fused_adagrad_flat": "pp pipil fff ppp",
    "dsb200_fused_lamb_flat": "ppp pipi pp l fffff ii ff pp p p",
    "dsb200_fused_lamb_scratch_floats": "l",
    "dsb200_multi_tensor_adam": "i ppppp i fffff iii ppp",
    "dsb200_norm_fwd": "pppppppp ii f ii p",
    "dsb200_norm_bwd_scratch_floats": "iii",
    "dsb200_norm_bwd": "pppppppppp ii iiii p",
    "dsb200_rms_norm_bwd_noalloc": "ppppppp iiiii p",
    "dsb200_rope_inplace": "ppp l iiiiiii p",
    "dsb200_gated_act_fwd": "ppp l iiii p",
    "dsb200_gated_act_bwd": "ppp l iiii p",
    "dsb200_bias_act": "ppp l iii p",
    "dsb200_cross_entropy_fwd": "pppp l i l l i p",
    "dsb200_cross_entropy_bwd": "pppp l i l l i p",
    "dsb200_add": "ppp l i p",
    "dsb200_sumsq": "p l i p p",
}
_RESTYPE_LONG = {"dsb200_norm_bwd_scratch_floats"}


def register_sigs(sigs, restype_long=()):
    """Later-loaded op modules add their entry points here."""
    _SIGS.update(sigs)
    _RESTYPE_LONG.update(restype_long)
    if _dev is not None:
        _apply_sigs(_dev)


def _apply_sigs(l):
    for name, sig in _SIGS.items():
        if hasattr(l, name):
            fn = getattr(l, name)
            fn.argtypes = [_T[c] for c in sig.replace(" ", "")]
            fn.restype = ctypes.c_longlong if name in _RESTYPE_LONG else ctypes.c_int


def lib():
    global _dev
    if _dev is None:
        with _lock:
            if _dev is None:
                l = _load("libdsb200.so", "build_device")
                l.dsb200_last_error.restype = ctypes.c_char_p
                _apply_sigs(l)
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
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return t.data_ptr() if t is not None else None


c_int = ctypes.c_int
c_float = ctypes.c_float
c_long = ctypes.c_int64
c_void_p = ctypes.c_void_p
