"""Key names and defaults of the `flops_profiler` config section (parity: deepspeed/profiling/constants.py).  The names are
generated from one table so a key and its default cannot drift apart: FLOPS_PROFILER_<KEY> is the key string and
FLOPS_PROFILER_<KEY>_DEFAULT its default value."""
FLOPS_PROFILER = "flops_profiler"

_SECTION_DEFAULTS = {
    "enabled": False,
    "recompute_fwd_factor": 0.0,
    "profile_step": 1,
    "module_depth": -1,
    "top_modules": 1,
    "detailed": True,
    "output_file": None,
}

for _key, _default in _SECTION_DEFAULTS.items():
    globals()[f"FLOPS_PROFILER_{_key.upper()}"] = _key
    globals()[f"FLOPS_PROFILER_{_key.upper()}_DEFAULT"] = _default
del _key, _default

FLOPS_PROFILER_FORMAT = "flops profiler should be enabled as:\n\"session_params\": {\n  \"flops_profiler\": {\n" + "".join(
    f"    \"{k}\": {'null' if v is None else str(v).lower() if isinstance(v, bool) else v},\n" for k, v in _SECTION_DEFAULTS.items()) + "  }\n}\n"
