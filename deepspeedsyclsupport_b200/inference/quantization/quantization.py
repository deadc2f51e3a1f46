"""Post-init group-wise weight quantization driven by the ds_config:

    "weight_quantization": {"post_init_quant": {"<substring of module name>": {"num_bits": 4, "group_size": 64,
                                                                                 "group_dim": 1, "symmetric": false}, ...}}

Every nn.Linear / nn.Embedding whose qualified name contains a key is replaced by its quantized drop-in (small weights
first, so the peak stays low).  ZeRO-3 partitioned weights are gathered one at a time for the conversion.
(reference: deepspeed/inference/quantization/quantization.py _init_group_wise_weight_quantization)"""
import gc

from ...utils.logging import logger
from . import layers
from .layers import QUANTIZATION_LAYER_MAPPINGS
from .utils import recursive_setattr


def _init_group_wise_weight_quantization(model, ds_config):
    assert "weight_quantization" in ds_config, "Please provide quantization config in ds_config"
    qcfg = ds_config["weight_quantization"]["post_init_quant"]
    zero3 = ds_config.get("zero_optimization", {}).get("stage", 0) == 3
    offload = "offload_param" in ds_config.get("zero_optimization", {})

    def size_of(m):
        w = m.weight
        return int(getattr(w, "ds_numel", w.numel()))

    mods = [(n, m) for n, m in model.named_modules() if type(m) in QUANTIZATION_LAYER_MAPPINGS]
    if not offload:
        mods.sort(key=lambda nm: size_of(nm[1]))
    by_key, count = {}, 0
    while mods:
        name, mod = mods.pop(0)
        hits = [k for k in qcfg if k in name]
        assert len(hits) <= 1, f"{name} matched multiple quantization key word {hits}"
        if not hits:
            continue
        key = hits[0]
        if zero3 and hasattr(mod.weight, "ds_id"):
            from ...runtime.zero import GatheredParameters
            with GatheredParameters([mod.weight]):
                new = QUANTIZATION_LAYER_MAPPINGS[type(mod)](qcfg[key], mod)
        else:
            new = QUANTIZATION_LAYER_MAPPINGS[type(mod)](qcfg[key], mod)
        recursive_setattr(model, name, new)
        by_key.setdefault(key, []).append(name)
        count += 1
        del mod
        gc.collect()
    layers.quantized_weight_registry.clear()
    logger.info(f"Group-wise weight quantization summary: convert {count} node(s) to quantized implementation")
    for key, names in by_key.items():
        logger.info(f"Key: {key}, matched modules:\n\t" + "\n\t".join(names))
    return model
