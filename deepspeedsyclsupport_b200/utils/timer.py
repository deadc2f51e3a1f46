"""Timing utilities.

* SynchronizedWallClockTimer : a registry of named stopwatches.  On a GPU a stopwatch records (start, end) CUDA-event
  pairs on the launching stream and only resolves them to milliseconds when somebody asks (`elapsed`, `mean`, `log`),
  so the training loop never synchronises because of timing; on hosts without a device it uses perf_counter.
* ThroughputTimer            : samples/s over optimizer steps, skipping the first `start_step` warm-up steps.
* NoopTimer                  : same surface, does nothing (wall_clock_breakdown off).
The public names and the timer-name constants follow the reference (deepspeed/utils/timer.py) because the engine,
the flops profiler and user scripts address timers by those names."""
import time

from ..accelerator import get_accelerator
from .logging import log_dist

FORWARD_MICRO_TIMER, FORWARD_GLOBAL_TIMER = "fwd_microstep", "fwd"
BACKWARD_MICRO_TIMER, BACKWARD_GLOBAL_TIMER = "bwd_microstep", "bwd"
BACKWARD_INNER_MICRO_TIMER, BACKWARD_INNER_GLOBAL_TIMER = "bwd_inner_microstep", "bwd_inner"
BACKWARD_REDUCE_MICRO_TIMER, BACKWARD_REDUCE_GLOBAL_TIMER = "bwd_allreduce_microstep", "bwd_allreduce"
STEP_MICRO_TIMER, STEP_GLOBAL_TIMER = "step_microstep", "step"

try:
    import psutil  # noqa: F401
    PSUTILS_INSTALLED = True
except ImportError:
    PSUTILS_INSTALLED = False


def trim_mean(data, trim_percent):
    """Mean of `data` after dropping the `trim_percent` smallest and largest fractions (robust to warm-up outliers)."""
    assert 0.0 <= trim_percent <= 1.0
    if len(data) == 0:
        return 0
    ordered = sorted(data)
    cut = int(round(len(ordered) * trim_percent))
    kept = ordered[cut:len(ordered) - cut] or ordered
    return sum(kept) / len(kept)


class _Stopwatch(object):
    """One named timer.  Intervals are kept as unresolved spans until read."""

    def __init__(self, name):
        self.name_ = name
        self.started_ = False
        self._on_host = get_accelerator().use_host_timers()
        self._open = None  # start marker of the running interval
        self._spans = []  # [(start, end)] events or perf_counter floats, not yet resolved
        self._resolved = []  # milliseconds of the intervals resolved since the last reset
        self.elapsed_records = None  # resolved milliseconds as of the last read

    def _mark(self):
        if self._on_host:
            return time.perf_counter()
        ev = get_accelerator().Event(enable_timing=True)
        ev.record()
        return ev

    def start(self):
        assert not self.started_, f"{self.name_} timer has already been started"
        self._open = self._mark()
        self.started_ = True

    def stop(self, reset=False, record=False):
        assert self.started_, "timer is not started"
        self._spans.append((self._open, self._mark()))
        self._open = None
        self.started_ = False

    def _resolve(self):
        if self._on_host:
            ms = [(b - a) * 1e3 for a, b in self._spans]
        else:
            ms = []
            for a, b in self._spans:
                b.synchronize()
                ms.append(a.elapsed_time(b))
        self._spans = []
        self._resolved += ms  # a read without reset keeps what was measured so far
        self.elapsed_records = list(self._resolved)
        return sum(self._resolved)

    def reset(self):
        self.started_ = False
        self._open = None
        self._spans = []
        self._resolved = []
        self.elapsed_records = None

    def elapsed(self, reset=True):
        """Total milliseconds since the last reset; a running timer is paused for the read and restarted."""
        was_running = self.started_
        if was_running:
            self.stop()
        total = self._resolve()
        if reset:
            self.reset()
        if was_running:
            self.start()
        return total

    def mean(self):
        self.elapsed(reset=False)
        return trim_mean(self.elapsed_records, 0.1)


class SynchronizedWallClockTimer(object):
    Timer = _Stopwatch

    def __init__(self):
        self.timers = {}

    def __call__(self, name):
        sw = self.timers.get(name)
        if sw is None:
            sw = self.timers[name] = _Stopwatch(name)
        return sw

    def get_timers(self):
        return self.timers

    @staticmethod
    def memory_usage():
        acc = get_accelerator()
        gb = 1024.0**3
        fields = (("mem_allocated", acc.memory_allocated()), ("max_mem_allocated", acc.max_memory_allocated()),
                  ("cache_allocated", acc.memory_cached()), ("max_cache_allocated", acc.max_memory_cached()))
        return "".join(" | {}: {:.4f} GB".format(k, v / gb) for k, v in fields)

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None):
        assert normalizer > 0.0
        parts = ["time (ms)"]
        for name in names:
            if name in self.timers:
                parts.append("{}: {:.2f}".format(name, self.timers[name].elapsed(reset=reset) / normalizer))
        line = " | ".join(parts)
        if memory_breakdown:
            line += self.memory_usage()
        log_dist(line, ranks=ranks or [0])

    def get_mean(self, names, normalizer=1.0, reset=True):
        assert normalizer > 0.0
        return {name: self.timers[name].mean() * 1000.0 / normalizer for name in names if name in self.timers}


class NoopTimer(object):

    class Timer(object):

        def start(self):
            pass

        def stop(self, **kwargs):
            pass

        def reset(self):
            pass

        def elapsed(self, **kwargs):
            return 0

        def mean(self):
            return 0

    def __init__(self):
        self.timer = self.Timer()

    def __call__(self, name):
        return self.timer

    def get_timers(self):
        return {}

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None):
        pass

    def get_mean(self, names, normalizer=1.0, reset=True):
        return {}


class ThroughputTimer(object):
    """start()/stop() around every micro-step; stop(global_step=True) on the optimizer-step boundary."""

    def __init__(self, batch_size, start_step=2, steps_per_output=50, monitor_memory=False, logging_fn=None):
        from .logging import logger
        self.batch_size = batch_size if batch_size is not None else 1
        self.start_step = start_step
        self.steps_per_output = steps_per_output
        self.monitor_memory = monitor_memory
        self.logging = logging_fn if logging_fn is not None else logger.info
        self.initialized = False
        self.started = False
        self.epoch_count = 0
        self.micro_step_count = 0
        self.global_step_count = 0
        self.total_elapsed_time = 0.0
        self.step_elapsed_time = 0.0
        self.start_time = 0.0
        self.end_time = 0.0

    def update_epoch_count(self):
        self.epoch_count += 1
        self.micro_step_count = 0

    def _warm(self):
        return self.global_step_count >= self.start_step

    def start(self):
        self.initialized = True
        self.started = True
        if self._warm():
            get_accelerator().synchronize()
            self.start_time = time.perf_counter()

    def stop(self, global_step=False, report_speed=True):
        if not self.started:
            return
        self.started = False
        self.micro_step_count += 1
        self.global_step_count += int(bool(global_step))
        if self.start_time <= 0:
            return
        get_accelerator().synchronize()
        self.end_time = time.perf_counter()
        dt = self.end_time - self.start_time
        self.total_elapsed_time += dt
        self.step_elapsed_time += dt
        if not global_step:
            return
        if report_speed and self.steps_per_output and self.global_step_count % self.steps_per_output == 0:
            acc = get_accelerator()
            self.logging(f"epoch={self.epoch_count}/micro_step={self.micro_step_count}/global_step={self.global_step_count}, "
                         f"RunningAvgSamplesPerSec={self.avg_samples_per_sec()}, "
                         f"CurrSamplesPerSec={self.batch_size / max(self.step_elapsed_time, 1e-12)}, "
                         f"MemAllocated={round(acc.memory_allocated() / 1024**3, 2)}GB, "
                         f"MaxMemAllocated={round(acc.max_memory_allocated() / 1024**3, 2)}GB")
        self.step_elapsed_time = 0.0

    def avg_samples_per_sec(self):
        if self.global_step_count <= 0:
            return float("-inf")
        timed_steps = max(self.global_step_count - self.start_step, 1)
        return self.batch_size / max(self.total_elapsed_time / timed_steps, 1e-12)
