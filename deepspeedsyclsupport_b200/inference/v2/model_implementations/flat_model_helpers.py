"""One contiguous buffer for all parameters of a TP rank (parity: inference/v2/model_implementations/flat_model_helpers.py:19-282).

`flatten_inference_model` copies every finished container parameter (and its auxiliary tensors: fp8 scales) into one byte
buffer at 256-byte aligned offsets, re-points the containers at views of that buffer, and returns the buffer with a JSON-able
description.  A serialised engine is that buffer + description per rank: loading it back is one file read, one host-to-device
copy and `restore_inference_model`, with no re-sharding or re-quantisation."""
from os import path
from typing import Dict, Iterable, List, Tuple

import torch

from ....runtime.config_utils import DeepSpeedConfigModel
from ..inference_parameter import STR_TO_DTYPE, InferenceParameter
from ..inference_utils import elem_size


def pad_to_aligned_offset(offset: int, alignment: int = 256) -> int:
    return ((offset + alignment - 1) // alignment) * alignment


class TensorMetadata(DeepSpeedConfigModel):
    dtype: str
    shape: List[int]
    strides: List[int]
    offset: int


class ParameterMetadata(DeepSpeedConfigModel):
    core_param: TensorMetadata
    aux_params: Dict[str, TensorMetadata] = {}


class LayerMetadata(DeepSpeedConfigModel):
    params: Dict[str, ParameterMetadata] = {}


class ModelMetadata(DeepSpeedConfigModel):
    policy: str = ""
    layers: Dict[str, LayerMetadata] = {}


def make_param_filename(base: str, rank: int, n_ranks: int) -> str:
    return path.join(base, f"params_rank_{rank}_of_{n_ranks}.pt")


def make_metadata_filename(base: str, rank: int, n_ranks: int) -> str:
    return path.join(base, f"metadata_rank_{rank}_of_{n_ranks}.json")


def make_model_config_filename(base: str) -> str:
    return path.join(base, "ds_model_config.pkl")


def _nbytes(t: torch.Tensor) -> int:
    return t.numel() * elem_size(t.dtype)


def _plan(t: torch.Tensor, offset: int) -> Tuple[TensorMetadata, int]:
    offset = pad_to_aligned_offset(offset)
    t = t.contiguous()
    return TensorMetadata(dtype=str(t.dtype), shape=list(t.shape), strides=list(t.stride()), offset=offset), offset + _nbytes(t)


def _view(buffer: torch.Tensor, md: TensorMetadata) -> torch.Tensor:
    dtype = STR_TO_DTYPE[md.dtype]
    n = 1
    for d in md.shape:
        n *= d
    raw = buffer[md.offset:md.offset + n * elem_size(dtype)]
    return raw.view(dtype).view(md.shape)


def flatten_inference_model(named_containers: Iterable, policy_name: str = "") -> Tuple[torch.Tensor, ModelMetadata]:
    """named_containers: iterable of (name, LayerContainer) -- `model.named_containers()`."""
    named_containers = list(named_containers)
    metadata, total, device = ModelMetadata(policy=policy_name), 0, None
    for lname, container in named_containers:
        layer = LayerMetadata()
        for pname, tensor in container.items():
            device = tensor.device if device is None else device
            core, total = _plan(tensor, total)
            pm = ParameterMetadata(core_param=core)
            for aname, aux in getattr(tensor, "aux_attrs", {}).items():
                pm.aux_params[aname], total = _plan(aux, total)
            layer.params[pname] = pm
        metadata.layers[lname] = layer
    buffer = torch.empty(pad_to_aligned_offset(total), dtype=torch.uint8, device=device or "cpu")
    for lname, container in named_containers:
        for pname, tensor in list(container.items()):
            pm = metadata.layers[lname].params[pname]
            core = _view(buffer, pm.core_param)
            core.copy_(torch.Tensor(tensor) if isinstance(tensor, InferenceParameter) else tensor)
            aux = {}
            for aname, amd in pm.aux_params.items():
                aux[aname] = _view(buffer, amd)
                aux[aname].copy_(tensor.aux_attrs[aname])
            container.direct_set(pname, InferenceParameter.initialize(core, **aux))  # the live parameter now aliases the buffer
    return buffer, metadata


def restore_inference_model(buffer: torch.Tensor, metadata: ModelMetadata, named_containers: Iterable) -> None:
    """Point every container parameter at its slice of `buffer` (already on the target device)."""
    if isinstance(metadata, dict):
        metadata = ModelMetadata(**metadata)
    containers = dict(named_containers)
    for lname, layer in metadata.layers.items():
        container = containers[lname]
        for pname, pm in layer.params.items():
            aux = {aname: _view(buffer, amd) for aname, amd in pm.aux_params.items()}
            container.direct_set(pname, InferenceParameter.initialize(_view(buffer, pm.core_param), **aux))
