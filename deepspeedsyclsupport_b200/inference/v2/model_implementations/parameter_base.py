"""Declarative model parameters (parity: inference/v2/model_implementations/parameter_base.py:14-255).

A parameter of the inference model is usually *derived* from several checkpoint tensors (q, k, v -> fused QKV; E expert matrices
-> one stacked tensor) and then sharded / re-laid-out for the kernels.  A `ParameterBase` subclass names its source tensors as
class annotations and implements `finalize()`:

    class UnfusedQKVParameter(ParameterBase):
        q_params: torch.Tensor
        k_params: torch.Tensor
        v_params: torch.Tensor

        def finalize(self):
            return self.inference_model.transform_qkv_param(torch.cat([self.q_params, self.k_params, self.v_params]))

Assigning the last missing source (`p.v_params = t`, or `p.experts[3] = t` for a `ParamList`) runs `finalize()` at once, so the
checkpoint tensors of a layer can be dropped as soon as the layer is complete instead of after the whole model is read."""
from typing import Callable, Dict, Optional

import torch


class ParamList:
    """Annotation of a list-valued source: `experts: ParamList("n_experts")` -- the length is the attribute of that name on the
    inference model (an int may be given directly)."""

    def __init__(self, count_attr):
        self.count_attr = count_attr

    def resolve(self, model) -> int:
        if isinstance(self.count_attr, int):
            return self.count_attr
        n = getattr(model, self.count_attr)
        return int(n() if callable(n) else n)


class _BoundList:
    """The list a parameter instance holds for one ParamList source."""

    def __init__(self, owner: "ParameterBase", n: int):
        self._owner, self._items = owner, [None] * n

    def __setitem__(self, idx: int, value: torch.Tensor) -> None:
        if self._items[idx] is not None:
            raise ValueError(f"source {idx} was already set")
        self._items[idx] = value
        self._owner._maybe_finalize()

    def __getitem__(self, idx):
        return self._items[idx]

    def __len__(self):
        return len(self._items)

    def __iter__(self):
        return iter(self._items)

    @property
    def complete(self) -> bool:
        return all(v is not None for v in self._items)

    def clear(self):
        self._items = [None] * len(self._items)


class ParameterBase:
    _tensor_sources: tuple = ()
    _list_sources: Dict[str, ParamList] = {}

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        tensors, lists = list(cls._tensor_sources), dict(cls._list_sources)
        for name, ann in cls.__dict__.get("__annotations__", {}).items():
            default = cls.__dict__.get(name)
            if isinstance(default, ParamList):  # `experts: ParamList = ParamList("n_experts")`
                lists[name] = default
            elif isinstance(ann, ParamList):
                lists[name] = ann
            elif ann is torch.Tensor or ann == "torch.Tensor":
                tensors.append(name)
        cls._tensor_sources, cls._list_sources = tuple(tensors), lists

    def __init__(self, model, on_complete: Optional[Callable[["ParameterBase"], None]] = None) -> None:
        object.__setattr__(self, "inference_model", model)
        object.__setattr__(self, "_on_complete", on_complete)
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_result", None)
        object.__setattr__(self, "_finalized", False)
        for name, spec in self._list_sources.items():
            object.__setattr__(self, name, _BoundList(self, spec.resolve(model)))

    # ---- sources
    @classmethod
    def sources(cls):
        return cls._tensor_sources + tuple(cls._list_sources)

    def __setattr__(self, name, value):
        if name in self._tensor_sources:
            if name in self._values:
                raise ValueError(f"{type(self).__name__}.{name} was already set")
            self._values[name] = value
            self._maybe_finalize()
        elif name in self._list_sources:
            raise AttributeError(f"{name} is a list source: assign its elements")
        else:
            object.__setattr__(self, name, value)

    def __getattr__(self, name):  # only reached when normal lookup fails
        values = object.__getattribute__(self, "_values")
        if name in values:
            return values[name]
        raise AttributeError(f"{type(self).__name__} has no attribute or unset source '{name}'")

    @property
    def touched(self) -> bool:
        """Whether any source has been provided."""
        return bool(self._values) or any(any(v is not None for v in getattr(self, n)) for n in self._list_sources)

    @property
    def complete(self) -> bool:
        return all(n in self._values for n in self._tensor_sources) and all(getattr(self, n).complete for n in self._list_sources)

    @property
    def is_finalized(self) -> bool:
        return self._finalized

    def _maybe_finalize(self):
        if self._finalized or not self.complete:
            return
        object.__setattr__(self, "_result", self.finalize())
        object.__setattr__(self, "_finalized", True)
        self._values.clear()  # the checkpoint tensors are no longer needed
        for n in self._list_sources:
            getattr(self, n).clear()
        if self._on_complete is not None:
            self._on_complete(self)

    @property
    def value(self):
        if not self._finalized:
            raise RuntimeError(f"{type(self).__name__} is missing sources: " + ", ".join(self.missing()))
        return self._result

    def missing(self):
        out = [n for n in self._tensor_sources if n not in self._values]
        for n in self._list_sources:
            out += [f"{n}[{i}]" for i, v in enumerate(getattr(self, n)) if v is None]
        return out

    def finalize(self) -> torch.Tensor:
        """Combine the sources and hand the result to the model's transform for this kind of parameter."""
        raise NotImplementedError()
