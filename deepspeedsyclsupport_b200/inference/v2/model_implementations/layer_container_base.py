"""Containers: the set of parameters of one transformer layer (or of the embedding / final-norm / unembedding part) plus the map
from checkpoint tensor names onto their sources (parity: inference/v2/model_implementations/layer_container_base.py:21-355).

    class LlamaTransformerContainer(LayerContainer):
        qkv_w: UnfusedQKVParameter
        attn_out_w: AttentionOutputParameter
        qkv_b: Optional[UnfusedQKVParameter]          # optional: absent from the checkpoint -> None
        PARAM_MAPPING = {
            "self_attn.q_proj.weight": "qkv_w.q_params",
            "mlp.experts.*.w1.weight": "experts_gu.gating_experts",   # '*' = index into a ParamList source
            "attn.c_attn.weight": ("qkv_w.params", "transpose"),      # (target, name of a transform method of the container)
            "lm_head.weight": ["unembed.params", "other.params"],     # one tensor feeding several parameters
        }

Finished parameters are read back as attributes (`container.qkv_w`) or dict-style (`container["qkv_w"]`, `.get`, `in`)."""
import re
import typing
from typing import Dict, Optional

import torch

from .parameter_base import ParameterBase


def _unwrap_optional(ann):
    """-> (parameter class, optional?)"""
    if typing.get_origin(ann) is typing.Union:
        args = [a for a in typing.get_args(ann) if a is not type(None)]
        if len(args) == 1:
            return args[0], True
    return ann, False


class LayerContainer:
    PARAM_MAPPING: Dict[str, object] = {}
    _param_types: Dict[str, type] = {}
    _optional: frozenset = frozenset()

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        types, optional = dict(cls._param_types), set(cls._optional)
        for name, ann in cls.__dict__.get("__annotations__", {}).items():
            ptype, opt = _unwrap_optional(ann)
            if isinstance(ptype, type) and issubclass(ptype, ParameterBase):
                types[name] = ptype
                if opt:
                    optional.add(name)
        cls._param_types, cls._optional = types, frozenset(optional)
        # compile the name patterns once per class; inherited mappings are kept unless overridden
        merged = {}
        for base in reversed(cls.__mro__):
            merged.update(base.__dict__.get("PARAM_MAPPING", {}))
        cls._compiled = []
        for pattern, target in merged.items():
            targets = target if isinstance(target, list) else [target]
            norm = []
            for t in targets:
                dest, transform = (t if isinstance(t, tuple) else (t, None))
                pname, _, source = dest.partition(".")
                if pname not in types:
                    raise ValueError(f"{cls.__name__}.PARAM_MAPPING['{pattern}'] targets unknown parameter '{pname}'")
                if source not in types[pname].sources():
                    raise ValueError(f"{cls.__name__}.PARAM_MAPPING['{pattern}']: {types[pname].__name__} has no source '{source}'")
                norm.append((pname, source, transform))
            rx = re.compile("^" + re.escape(pattern).replace(r"\*", r"(\d+)") + "$")
            cls._compiled.append((rx, norm))

    def __init__(self, model) -> None:
        self.inference_model = model
        self._params = {name: ptype(model) for name, ptype in self._param_types.items()}
        self._finalized_optional = False

    # ---- loading
    @classmethod
    def mapped_names(cls):
        return [rx.pattern for rx, _ in cls._compiled]

    def set_dependency(self, dep_name: str, dep_value: torch.Tensor) -> bool:
        """Route one checkpoint tensor (name relative to this container).  Returns False when no pattern matches."""
        hit = False
        for rx, targets in self._compiled:
            m = rx.match(dep_name)
            if not m:
                continue
            hit = True
            for pname, source, transform in targets:
                value = getattr(self, transform)(dep_value) if transform else dep_value
                param = self._params[pname]
                if source in param._list_sources:
                    getattr(param, source)[int(m.group(1))] = value
                else:
                    setattr(param, source, value)
        return hit

    def direct_set(self, name: str, value) -> None:
        """Install an already transformed tensor (restoring a flattened model)."""
        p = self._params[name]
        object.__setattr__(p, "_result", value)
        object.__setattr__(p, "_finalized", True)

    # ---- state
    @property
    def is_populated(self) -> bool:
        """Every required parameter is complete and no optional one is half-filled."""
        for name, p in self._params.items():
            if p.is_finalized:
                continue
            if name in self._optional and not p.touched:
                continue
            return False
        return True

    def missing(self):
        out = []
        for name, p in self._params.items():
            if not p.is_finalized and not (name in self._optional and not p.touched):
                out += [f"{name}.{m}" for m in p.missing()]
        return out

    # ---- access
    def _value(self, name):
        p = self._params[name]
        if p.is_finalized:
            return p.value
        if name in self._optional and not p.touched:
            return None
        raise RuntimeError(f"{type(self).__name__}.{name} is incomplete: missing {p.missing()}")

    def __getattr__(self, name):
        params = self.__dict__.get("_params")
        if params is not None and name in params:
            return self._value(name)
        raise AttributeError(name)

    def __getitem__(self, name):
        v = self._value(name)
        if v is None:
            raise KeyError(name)
        return v

    def get(self, name, default=None):
        if name not in self._params:
            return default
        v = self._value(name)
        return default if v is None else v

    def __contains__(self, name):
        return name in self._params and self._value(name) is not None

    def items(self):
        """(name, tensor) of every present parameter, in declaration order."""
        for name in self._params:
            v = self._value(name)
            if v is not None:
                yield name, v

    def param_names(self):
        return list(self._params)
