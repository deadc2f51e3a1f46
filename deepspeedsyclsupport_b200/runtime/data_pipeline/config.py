"""data_efficiency config (parity: deepspeed/runtime/data_pipeline/config.py)."""
import copy

DATA_EFFICIENCY = "data_efficiency"


def get_data_efficiency_config(param_dict):
    sub = copy.deepcopy(param_dict.get(DATA_EFFICIENCY, {}))
    out = dict(enabled=sub.get("enabled", False), seed=sub.get("seed", 1234))
    ds = sub.get("data_sampling", {})
    cl = ds.get("curriculum_learning", {})
    out["data_sampling"] = dict(enabled=ds.get("enabled", False), num_epochs=ds.get("num_epochs", 1000),
                                num_workers=ds.get("num_workers", 0),
                                curriculum_learning=dict(enabled=cl.get("enabled", False), **{k: v for k, v in cl.items() if k != "enabled"}))
    dr = sub.get("data_routing", {})
    ltd = dr.get("random_ltd", {})
    out["data_routing"] = dict(enabled=dr.get("enabled", False),
                               random_ltd=dict(enabled=ltd.get("enabled", False), **{k: v for k, v in ltd.items() if k != "enabled"}))
    return out
