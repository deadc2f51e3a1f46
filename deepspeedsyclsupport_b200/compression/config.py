"""compression_training config (parity: deepspeed/compression/config.py). Kept as a validated nested dict."""
import copy

COMPRESSION_TRAINING = "compression_training"
SHARED_PARAMETERS = "shared_parameters"
DIFFERENT_GROUPS = "different_groups"
TECHNIQUES = ["weight_quantization", "activation_quantization", "sparse_pruning", "row_pruning", "head_pruning",
              "channel_pruning"]
LAYER_REDUCTION = "layer_reduction"

_SHARED_DEFAULTS = {
    "weight_quantization": dict(enabled=False, quantizer_kernel=False, schedule_offset=0, quantize_groups=1,
                                quantize_verbose=False, quantization_type="symmetric", quantize_weight_in_forward=False,
                                rounding="nearest", fp16_mixed_quantize=dict(enabled=False, quantize_change_ratio=0.001)),
    "activation_quantization": dict(enabled=False, quantization_type="symmetric", range_calibration="dynamic", schedule_offset=1000),
    "sparse_pruning": dict(enabled=False, method="l1", schedule_offset=1000, schedule_offset_end=1000, schedule_offset_stride=1,
                           block_pattern="4x1", dense_ratio=0.1, excluded_modules=[]),
    "row_pruning": dict(enabled=False, method="l1", schedule_offset=1000),
    "head_pruning": dict(enabled=False, method="topk", schedule_offset=1000, num_heads=None),
    "channel_pruning": dict(enabled=False, method="l1", schedule_offset=1000),
}


def get_compression_config(param_dict):
    sub = copy.deepcopy(param_dict.get(COMPRESSION_TRAINING, {}))
    out = {}
    for tech in TECHNIQUES:
        t = sub.get(tech, {})
        shared = dict(_SHARED_DEFAULTS[tech])
        shared.update(t.get(SHARED_PARAMETERS, {}))
        groups = t.get(DIFFERENT_GROUPS, {})
        for name, g in groups.items():
            assert "params" in g, f"compression group {name} of {tech} needs 'params'"
            g.setdefault("modules", ["*"])
            g.setdefault("related_modules", None)
        if tech == "head_pruning" and shared["enabled"]:
            assert shared.get("num_heads") is not None, "head_pruning requires num_heads"
        out[tech] = {SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: groups}
    lr = sub.get(LAYER_REDUCTION, {})
    out[LAYER_REDUCTION] = dict(enabled=lr.get("enabled", False), keep_number_layer=lr.get("keep_number_layer"),
                                module_name_prefix=lr.get("module_name_prefix", ""), teacher_layer=lr.get("teacher_layer", []),
                                other_module_name=lr.get("other_module_name", []))
    return out


def get_quantize_enabled(param_dict):
    return param_dict.get("quantize_training", {}).get("enabled", False)
