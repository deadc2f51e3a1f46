"""`ds_report`: what is installed and whether the native ops can run here.

Two tables: (1) one row per op name with "installed" (its library -- libdsb200.so for CUDA kernels, libdsb200_host.so for
host natives -- exists in-tree) and "compatible" (the builder's own check); (2) general environment facts.  Everything is
built ahead of time for sm_100a, so unlike the reference's report (deepspeed/env_report.py:30-194) there is no JIT column.
"""
import argparse
import os
import subprocess

import torch

from .ops.op_builder import ALL_OPS
from .ops.op_builder import builder as B
from .version import __version__

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"


def _paint(colour, text):
    return f"{colour}{text}{END}"


SUCCESS, OKAY, WARNING, FAIL, INFO = _paint(GREEN, " [SUCCESS] "), _paint(GREEN, "[OKAY]"), _paint(YELLOW, "[WARNING]"), _paint(RED, "[FAIL]"), "[INFO]"
okay, warning = OKAY, WARNING
color_len = len(GREEN) + len(END)


def _dotted(cells, widths):
    """'name ...... value ..... value' with each column padded by dots to its width (colour codes do not count)."""
    out = []
    for cell, width in zip(cells, widths):
        visible = len(cell) - (color_len if "\033[" in cell else 0)
        out.append(cell + (" " + "." * max(1, width - visible) if width else ""))
    return " ".join(out)


def op_report(verbose=True):
    widths = (24, 12, 0)
    rule = "-" * 64
    have = {False: os.path.exists(os.path.join(B.LIB_DIR, "libdsb200.so")), True: os.path.exists(os.path.join(B.LIB_DIR, "libdsb200_host.so"))}
    print(rule)
    print("DeepSpeed C++/CUDA extension op report")
    print(rule)
    print("NOTE: every op resolves to one of two in-tree libraries built ahead of time for sm_100a\n"
          "      (libdsb200.so: CUDA kernels, libdsb200_host.so: host natives); there is no JIT matrix.")
    print(rule)
    print(_dotted(("op name", "installed", "compatible"), widths))
    print(rule)
    yes, no = _paint(GREEN, "[YES]"), _paint(YELLOW, "[NO]")
    for op_name, builder in ALL_OPS.items():
        print(_dotted((op_name, yes if have[bool(builder.HOST)] else no, OKAY if builder.is_compatible(verbose) else no), widths))
    print(rule)


def nvcc_version():
    nvcc = B.nvcc_path()
    if nvcc is None:
        return _paint(RED, " [FAIL] cannot find nvcc ")
    try:
        words = subprocess.check_output([nvcc, "-V"], universal_newlines=True).split()
    except (FileNotFoundError, subprocess.CalledProcessError):
        return _paint(RED, " [FAIL] nvcc missing ")
    return words[words.index("release") + 1].rstrip(",")


def human_readable_size(size):
    for unit in ("B", "KB", "MB", "GB"):
        if size < 1024:
            return f"{size:.2f} {unit}"
        size /= 1024
    return f"{size:.2f} TB"


def get_shm_size():
    """(human-readable size of /dev/shm, list of warnings) -- NCCL and the SHM all-reduce need a roomy /dev/shm in containers."""
    try:
        st = os.statvfs("/dev/shm")
    except (OSError, ValueError):
        return "UNKNOWN", None
    total = st.f_frsize * st.f_blocks
    notes = []
    if total < 512 << 20:
        notes.append(_paint(YELLOW, " [WARNING] /dev/shm size might be too small, if running in docker increase to at least --shm-size='1gb' "))
    return human_readable_size(total), notes


def debug_report():
    cuda = torch.cuda.is_available()
    facts = [("torch install path", torch.__path__), ("torch version", torch.__version__),
             ("deepspeed install path", [os.path.dirname(__file__)]), ("deepspeed info", f"{__version__}, B200-native (sm_100a)"),
             ("torch cuda version", torch.version.cuda), ("nvcc version", nvcc_version()), ("cuda available", cuda),
             ("device", torch.cuda.get_device_name(0) if cuda else "none"), ("shared memory (/dev/shm) size", get_shm_size()[0])]
    print("DeepSpeed general environment info:")
    for name, value in facts:
        print(_dotted((name, str(value)), (34, 0)))


def parse_arguments():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hide_operator_status", action="store_true", help="skip the per-op installed / compatible table")
    ap.add_argument("--hide_errors_and_warnings", action="store_true", help="do not print builder warnings")
    return ap.parse_args()


def main(hide_operator_status=False, hide_errors_and_warnings=False):
    if not hide_operator_status:
        op_report(verbose=not hide_errors_and_warnings)
    debug_report()


def cli_main():
    a = parse_arguments()
    main(hide_operator_status=a.hide_operator_status, hide_errors_and_warnings=a.hide_errors_and_warnings)


if __name__ == "__main__":
    main()
