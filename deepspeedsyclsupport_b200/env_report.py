"""`ds_report`: environment + op compatibility report (parity: deepspeed/env_report.py:30-194)."""
import argparse
import os
import shutil
import subprocess

import torch

from .ops.op_builder import ALL_OPS
from .ops.op_builder import builder as B
from .version import __version__

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
SUCCESS = f"{GREEN} [SUCCESS] {END}"
OKAY = f"{GREEN}[OKAY]{END}"
WARNING = f"{YELLOW}[WARNING]{END}"
FAIL = f"{RED}[FAIL]{END}"
INFO = "[INFO]"
color_len = len(GREEN) + len(END)
okay = f"{GREEN}[OKAY]{END}"
warning = f"{YELLOW}[WARNING]{END}"


def op_report(verbose=True):
    max_dots = 23
    max_dots2 = 11
    h = ["op name", "installed", "compatible"]
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    print("DeepSpeed C++/CUDA extension op report")
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    print("NOTE: every op resolves to one of two in-tree libraries built ahead of time for sm_100a\n"
          "      (libdsb200.so: CUDA kernels, libdsb200_host.so: host natives); there is no JIT matrix.")
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    print(h[0], "." * (max_dots - len(h[0])), h[1], "." * (max_dots2 - len(h[1])), h[2])
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    dev_ok = os.path.exists(os.path.join(B.LIB_DIR, "libdsb200.so"))
    host_ok = os.path.exists(os.path.join(B.LIB_DIR, "libdsb200_host.so"))
    installed = f"{GREEN}[YES]{END}"
    no = f"{YELLOW}[NO]{END}"
    for op_name, builder in ALL_OPS.items():
        dots = "." * (max_dots - len(op_name))
        is_installed = installed if (host_ok if builder.HOST else dev_ok) else no
        is_compatible = OKAY if builder.is_compatible(verbose) else no
        dots2 = "." * ((len(h[1]) + (max_dots2 - len(h[1]))) - (len(is_installed) - color_len))
        print(op_name, dots, is_installed, dots2, is_compatible)
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))


def nvcc_version():
    nvcc = B.nvcc_path()
    if nvcc is None:
        return f"{RED} [FAIL] cannot find nvcc {END}"
    try:
        output = subprocess.check_output([nvcc, "-V"], universal_newlines=True)
    except FileNotFoundError:
        return f"{RED} [FAIL] nvcc missing {END}"
    output_split = output.split()
    release_idx = output_split.index("release")
    release = output_split[release_idx + 1].replace(",", "").split(".")
    return ".".join(release)


def get_shm_size():
    try:
        shm_stats = os.statvfs("/dev/shm")
    except (OSError, FileNotFoundError, ValueError):
        return "UNKNOWN", None
    shm_size = shm_stats.f_frsize * shm_stats.f_blocks
    shm_hbytes = human_readable_size(shm_size)
    warn = []
    if shm_size < 512 * 1024 ** 2:
        warn.append(f" {YELLOW} [WARNING] /dev/shm size might be too small, if running in docker increase to at least --shm-size='1gb' {END}")
    return shm_hbytes, warn


def human_readable_size(size):
    units = ["B", "KB", "MB", "GB", "TB"]
    i = 0
    while size >= 1024:
        size /= 1024
        i += 1
    return f"{size:.2f} {units[i]}"


def debug_report():
    max_dots = 33
    report = [("torch install path", torch.__path__), ("torch version", torch.__version__), ("deepspeed install path", [os.path.dirname(__file__)]),
              ("deepspeed info", f"{__version__}, B200-native (sm_100a)"), ("torch cuda version", torch.version.cuda),
              ("nvcc version", nvcc_version()), ("cuda available", torch.cuda.is_available()),
              ("device", torch.cuda.get_device_name(0) if torch.cuda.is_available() else "none"),
              ("shared memory (/dev/shm) size", get_shm_size()[0])]
    print("DeepSpeed general environment info:")
    for name, value in report:
        print(name, "." * (max_dots - len(name)), value)


def parse_arguments():
    parser = argparse.ArgumentParser()
    parser.add_argument("--hide_operator_status", action="store_true", help="Suppress display of installation and compatibility statuses of DeepSpeed operators. ")
    parser.add_argument("--hide_errors_and_warnings", action="store_true", help="Suppress warning and error messages.")
    return parser.parse_args()


def main(hide_operator_status=False, hide_errors_and_warnings=False):
    if not hide_operator_status:
        op_report(verbose=not hide_errors_and_warnings)
    debug_report()


def cli_main():
    args = parse_arguments()
    main(hide_operator_status=args.hide_operator_status, hide_errors_and_warnings=args.hide_errors_and_warnings)


if __name__ == "__main__":
    main()
