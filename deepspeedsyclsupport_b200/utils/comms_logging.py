"""Per-op communication log: latency, message size, algbw / busbw with the nccl-tests formulas and an optional
straggler report.  Parity: deepspeed/utils/comms_logging.py:34-178."""
import math

from .logging import log_dist


def get_caller_func(frame=3):
    import sys
    return sys._getframe(frame).f_code.co_name


def convert_size(size_bytes):
    if size_bytes == 0:
        return "0B"
    names = ("B", "KB", "MB", "GB", "TB", "PB")
    i = int(math.floor(math.log(size_bytes, 1024)))
    return "%s %s" % (round(size_bytes / math.pow(1024, i), 2), names[i])


def calc_bw_log(comm_op, size, duration_ms, n):
    """Returns (algbw, busbw) in Gbps, same correction factors as nccl-tests."""
    duration = max(duration_ms / 1e3, 1e-12)
    if comm_op in ("all_to_all_single", "all_to_all"):
        tput, busbw = size / duration, (size / duration) * ((n - 1) / n)
    elif comm_op in ("all_gather", "all_gather_into_tensor", "reduce_scatter", "reduce_scatter_tensor",
                     "all_gather_base", "reduce_scatter_base"):
        size *= n
        tput, busbw = size / duration, (size / duration) * ((n - 1) / n)
    elif comm_op in ("all_reduce", "inference_all_reduce", "all_reduce_coalesced"):
        tput, busbw = size * 2 / duration, (size / duration) * (2 * (n - 1) / n)
    else:  # send/recv/broadcast/reduce/barrier...
        tput = busbw = size / duration
    return tput * 8 / 1e9, busbw * 8 / 1e9


class CommsLogger:

    def __init__(self):
        self.comms_dict = {}
        self.verbose = False
        self.debug = False
        self.prof_ops = []
        self.prof_all = True
        self.enabled = False

    def configure(self, comms_config):
        self.enabled = comms_config.comms_logger_enabled
        if self.enabled:
            c = comms_config.comms_logger
            self.verbose, self.debug, self.prof_ops, self.prof_all = c.verbose, c.debug, c.prof_ops, c.prof_all

    def start_profiling_comms(self):
        self.prof_all = True

    def stop_profiling_comms(self):
        self.prof_all = False

    def start_profiling_op(self, op_name_list):
        self.prof_ops = list(set(self.prof_ops) | set(op_name_list))

    def stop_profiling_op(self, op_name_list):
        self.prof_ops = [op for op in self.prof_ops if op not in op_name_list]

    def append(self, raw_name, record_name, latency, msg_size, world_size):
        algbw, busbw = calc_bw_log(raw_name, msg_size, latency, world_size)
        per_op = self.comms_dict.setdefault(record_name, {})
        if msg_size in per_op:
            per_op[msg_size][0] += 1
            per_op[msg_size][1].append(latency)
            per_op[msg_size][2].append(algbw)
            per_op[msg_size][3].append(busbw)
        else:
            per_op[msg_size] = [1, [latency], [algbw], [busbw]]
        if self.verbose:
            log_dist(f"comm op: {record_name} | time (ms): {latency:.2f} | msg size: {convert_size(msg_size)} | "
                     f"algbw (Gbps): {algbw:.2f} | busbw (Gbps): {busbw:.2f}", ranks=[0])

    def log_all(self, print_log=True, show_straggler=False):
        from ..utils.timer import trim_mean
        lines = [f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <20}{'Total Latency(ms)': <20}"
                 f"{'Avg Latency(ms)': <20}{'tput_avg (Gbps)': <20}{'busbw_avg (Gbps)': <20}"]
        for record_name, sizes in self.comms_dict.items():
            lines.append(record_name)
            for msg_size, vals in sorted(sizes.items()):
                count, lats, algs, buss = vals
                lines.append(f"{' ': <20}{convert_size(msg_size): <20}{count: <20}{sum(lats): <20.2f}"
                             f"{trim_mean(lats, 0.1): <20.2f}{trim_mean(algs, 0.1): <20.2f}{trim_mean(buss, 0.1): <20.2f}")
        if show_straggler:
            import torch
            from .. import comm as dist
            lines.append("_______________________________")
            lines.append("Breakdown with straggler effect")
            lines.append(f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <20}{'Total comm lat(ms)': <20}"
                         f"{'Total straggler(ms)': <20}{'Avg comm lat(ms)': <20}{'Avg straggler(ms)': <20}")
            for record_name, sizes in self.comms_dict.items():
                lines.append(record_name)
                for msg_size, vals in sorted(sizes.items()):
                    count = vals[0]
                    lats = torch.tensor(vals[1])
                    min_lats = lats.clone()
                    if dist.is_initialized():
                        dev = "cuda" if torch.cuda.is_available() else "cpu"
                        min_lats = min_lats.to(dev)
                        dist.all_reduce(min_lats, op=dist.ReduceOp.MIN)
                        min_lats = min_lats.cpu()
                    total_lat = min_lats.sum().item()
                    total_strag = (lats - min_lats).sum().item()
                    lines.append(f"{' ': <20}{convert_size(msg_size): <20}{count: <20}{total_lat: <20.2f}"
                                 f"{total_strag: <20.2f}{total_lat / count: <20.2f}{total_strag / count: <20.2f}")
        out = "\n".join(lines)
        if print_log:
            log_dist("\n" + out, ranks=[0])
        return out
