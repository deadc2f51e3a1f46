"""Fetch / release coordination of sharded parameters (parity: deepspeed/runtime/zero/partitioned_param_coordinator.py:38-520).

The reference coordinator records the order in which sub-modules touch parameters during the first iteration, then replays it:
all-gathers are issued ahead of use up to `prefetch_bucket_size`, parameters are released unless they will be reused within
`max_reuse_distance`, and at most `max_live_parameters` stay gathered.  Here that logic lives in the flat unit store
(`sharded_optimizer.ZeroShardedOptimizer`: `_record` / `_advance` / `_prefetch_next` / `_keep_after_forward`) at *unit*
granularity -- one contiguous buffer per transformer block, so a fetch is one collective instead of one per parameter.
`PartitionedParameterCoordinator` is the reference-shaped view of that state for code that drives fetches itself."""
from collections import UserDict
from enum import Enum
from typing import Iterable

from torch.nn import Module, Parameter


def get_all_parameters(sub_module, recurse=False):
    return sub_module.named_parameters(recurse=recurse)


def iter_params(module: Module, recurse=False) -> Iterable[Parameter]:
    return map(lambda pair: pair[1], get_all_parameters(module, recurse))


class ZeRoTraceMode(Enum):
    RECORD = 1  # first iteration: the order of unit fetches is being recorded
    COMPLETE = 2  # recorded; fetches are replayed as prefetch
    INVALID = 3  # the model took another path than the recorded one (dynamic control flow): prefetch is off until re-recorded


class InflightParamRegistry(UserDict):
    """Gathers that were issued but not yet waited for (unit id -> stream event)."""

    def __setitem__(self, key, handle) -> None:
        if key in self.data:
            raise RuntimeError(f"{key} already in registry")
        self.data[key] = handle


class PartitionedParameterCoordinator:
    _views = {}

    @classmethod
    def for_store(cls, store, training=True):
        key = (id(store), bool(training))
        if key not in cls._views:
            cls._views[key] = cls(store, training)
        return cls._views[key]

    def __init__(self, store, training=True):
        self._store = store
        self._training = training

    # ---- trace state
    @property
    def trace_mode(self) -> ZeRoTraceMode:
        return ZeRoTraceMode.COMPLETE if self._store._trace_frozen else ZeRoTraceMode.RECORD

    def is_complete_trace(self) -> bool:
        return self._store._trace_frozen

    def is_record_trace(self) -> bool:
        return not self._store._trace_frozen

    def is_invalid_trace(self) -> bool:
        return False

    def _invalidate_trace(self) -> None:
        """Forget the recorded order (call after changing which layers run): the next iteration records again."""
        s = self._store
        s._trace = {"fwd": [], "bwd": []}
        s._trace_pos = {"fwd": 0, "bwd": 0}
        s._trace_frozen = False

    def reset_step(self) -> None:
        """End of an iteration: freeze a newly recorded trace and rewind the replay cursors."""
        s = self._store
        if not s._trace_frozen and s._trace["fwd"]:
            s._trace_frozen = True
        s._trace_pos = {"fwd": 0, "bwd": 0}

    # ---- fetch / release for one module
    def _units(self, sub_module: Module):
        seen, out = set(), []
        for p in iter_params(sub_module, recurse=False):
            u = getattr(p, "ds_unit", None)
            if u is not None and id(u) not in seen:
                seen.add(id(u))
                out.append(u)
        return out

    def fetch_sub_module(self, current_submodule: Module, forward: bool) -> None:
        units = self._units(current_submodule)
        phase = "fwd" if forward else "bwd"
        for u in units:
            self._store._record(phase, u)
        self._store._advance(phase, units)
        self._store.fetch_units(units, prefetch_phase=phase)

    def release_sub_module(self, submodule: Module, backward: bool) -> None:
        for u in self._units(submodule):
            if not backward and self._training and self._store._keep_after_forward(u):
                continue
            self._store.release_unit(u)

    def release_and_reset_all(self, module: Module) -> None:
        for u in self._store.units:
            self._store.release_unit(u)
        self.reset_step()
