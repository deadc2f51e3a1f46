"""Version / build info (parity: deepspeed/git_version_info.py): version string, git hash/branch when available, and the
installed / compatible op tables used by ds_report."""
import os
import subprocess

from .version import __version__ as version


def _git(*args):
    try:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        return subprocess.check_output(["git", *args], cwd=root, stderr=subprocess.DEVNULL).decode().strip()
    except Exception:
        return "unknown"


git_hash = _git("rev-parse", "--short", "HEAD")
git_branch = _git("rev-parse", "--abbrev-ref", "HEAD")


def _op_tables():
    from .ops.op_builder import ALL_OPS
    from .ops.op_builder import builder as B
    dev_ok = os.path.exists(os.path.join(B.LIB_DIR, "libdsb200.so"))
    host_ok = os.path.exists(os.path.join(B.LIB_DIR, "libdsb200_host.so"))
    installed = {name: (host_ok if b.HOST else dev_ok) for name, b in ALL_OPS.items()}
    compatible = {name: bool(b.is_compatible(False)) for name, b in ALL_OPS.items()}
    return installed, compatible


try:
    installed_ops, compatible_ops = _op_tables()
except Exception:
    installed_ops, compatible_ops = {}, {}
try:
    import torch
    torch_info = {"version": torch.__version__, "cuda_version": torch.version.cuda, "hip_version": getattr(torch.version, "hip", None)}
except Exception:
    torch_info = {"version": "0.0", "cuda_version": "0.0", "hip_version": "0.0"}
