"""Engine construction: checkpoint -> family policy -> containers -> model -> InferenceEngineV2.
Parity: inference/v2/engine_factory.py:28-123 (build_hf_engine, build_engine_from_ds_checkpoint).  The family policies
(model_implementations/<family>/policy.py) own the HF tensor-name mapping and the config translation."""
import json
import os
import pickle

import torch

from .checkpoint import HuggingFaceCheckpointEngine, InMemoryModelEngine
from .config_v2 import RaggedInferenceEngineConfig
from .engine_v2 import InferenceEngineV2
from .model_implementations import SPEC_FROM_HF, ArchSpec, RaggedTransformer, policy_for  # noqa: F401


def build_model_from_checkpoint(model_type, hf_config: dict, checkpoint_engine, engine_config, tp_group=None, device=None, dtype=torch.bfloat16):
    """One TP rank's populated model of the family serving `model_type`."""
    policy = policy_for(model_type)(dict(hf_config, model_type=model_type), checkpoint_engine=checkpoint_engine)
    return policy.build_model(engine_config, tp_group, device=device, dtype=dtype)


def build_engine_from_state_dict(model_type, hf_config: dict, state_dict: dict, engine_config: RaggedInferenceEngineConfig,
                                 dtype=torch.bfloat16):
    """Engine from an in-memory HF-named state dict (used for random-init models when no checkpoint is on disk)."""
    policy_for(model_type)  # unsupported families fail here, before any process group is built

    def builder(tp_group, device):
        return build_model_from_checkpoint(model_type, hf_config, InMemoryModelEngine(state_dict), engine_config, tp_group, device, dtype)

    return InferenceEngineV2(builder, engine_config)


def build_hf_engine(path: str, engine_config: RaggedInferenceEngineConfig, debug_level: int = 20):
    if os.path.exists(os.path.join(path, "ds_model_config.pkl")):
        return build_engine_from_ds_checkpoint(path, engine_config, debug_level=debug_level)
    ck = HuggingFaceCheckpointEngine(path)
    model_type = ck.model_config["model_type"]
    policy_for(model_type)
    dtype = {"float16": torch.float16, "float32": torch.float32}.get(str(ck.model_config.get("torch_dtype", "bfloat16")), torch.bfloat16)

    def builder(tp_group, device):
        # the checkpoint is streamed: every tensor is sharded and moved as soon as its parameter is complete
        return build_model_from_checkpoint(model_type, ck.model_config, ck, engine_config, tp_group, device, dtype)

    return InferenceEngineV2(builder, engine_config)


def build_engine_from_ds_checkpoint(path: str, engine_config: RaggedInferenceEngineConfig, debug_level: int = 20):
    """Rebuild from `InferenceEngineV2.serialize()` output: per rank one flat byte buffer (already TP-sharded and laid out for the
    kernels, fp8 scales included) + its description; the containers are pointed at slices of the buffer."""
    from .inference_parameter import STR_TO_DTYPE
    from .model_implementations.flat_model_helpers import (make_metadata_filename, make_model_config_filename, make_param_filename,
                                                           restore_inference_model)
    with open(make_model_config_filename(path), "rb") as f:
        saved = pickle.load(f)

    def builder(tp_group, device):
        m = RaggedTransformer(saved["spec"], None, tp_group=tp_group, dtype=STR_TO_DTYPE[saved["dtype"]], device=device,
                              quantization_mode=saved.get("quantization_mode", ""), engine_config=engine_config)
        buffer = torch.load(make_param_filename(path, m.tp_rank, m.tp_size), map_location="cpu").to(device)
        with open(make_metadata_filename(path, m.tp_rank, m.tp_size)) as f:
            restore_inference_model(buffer, json.load(f), m.named_containers())
        m.flattened_params = buffer
        m.check_populated()
        return m

    return InferenceEngineV2(builder, engine_config)
