"""hp/lp linkage API names (parity: deepspeed/utils/mixed_precision_linkage.py).  In the flat unit store every low
precision parameter is a view at a known offset of its unit, and the fp32 master / optimizer state live at the same
offsets of the flat shard; the accessors in utils/tensor_fragment.py resolve that mapping, so linking is a no-op that
only validates the parameters are managed."""
from .tensor_fragment import safe_get_full_fp32_param, safe_get_full_grad, safe_get_full_optimizer_state  # noqa: F401


def link_hp_params(lp_param_list, flat_hp_partition, gradient_dict, offload_gradient_dict, use_offload, param_group_index, partition_start,
                   partition_size, partition_optimizer_state, dp_group):
    for p in lp_param_list:
        assert hasattr(p, "ds_unit"), "parameter is not managed by the flat unit store"


def lazy_init_hp_params_optimizer_state(lp_param_list, flat_hp_partition, optimizer_state):
    return
