"""Counts and times parameter-movement events of ZeRO stage 3 -- fetches, releases, gathers -- by event name (parity:
deepspeed/runtime/zero/partitioned_param_profiler.py:10-63).  A profiler without timers is inert, so call sites need no
`if profiling:` guards."""
from ...utils.logging import log_dist


class _Event:
    __slots__ = ("name", "count", "num_elem")

    def __init__(self, name):
        self.name, self.count, self.num_elem = name, 0, 0

    def reset(self):
        self.count = self.num_elem = 0

    def increment(self, numel):
        self.count += 1
        self.num_elem += numel

    def __repr__(self):
        return f"{self.name}: count = {self.count}, numel = {self.num_elem}"


class PartitionedParameterProfiler(object):
    EventCounter = _Event

    def __init__(self, timers):
        self.timers = timers
        self.event_counters = {}

    @property
    def _active(self):
        return self.timers is not None

    def reset_events(self):
        for event in self.event_counters.values():
            event.reset()

    def start_event(self, name):
        if self._active:
            self.event_counters.setdefault(name, _Event(name))
            self.timers(name).start()

    def stop_event(self, name, num_elem):
        if self._active:
            if name not in self.event_counters:
                raise AssertionError(f"stop_event('{name}') without a matching start_event")
            self.event_counters[name].increment(num_elem)
            self.timers(name).stop()

    def log_events(self):
        for event in self.event_counters.values():
            log_dist(repr(event), ranks=[0])
        if self._active and self.event_counters:
            self.timers.log(names=list(self.event_counters))
