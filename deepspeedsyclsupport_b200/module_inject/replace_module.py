"""Module replacement driver (parity: deepspeed/module_inject/replace_module.py:182-498): walks the model, applies the
AutoTP policy (or a user injection policy {BlockClass: ("linear.names", ...)}) to every matching block."""
import torch
from torch import nn

from .. import comm as dist
from .auto_tp import AutoTP, ReplaceWithTensorSlicing  # noqa: F401


class GroupQuantizer:
    """Group-wise int8 weight quantization used when loading quantized inference checkpoints."""

    def __init__(self, q_int8=True, group_size=1, num_bits=8, num_groups=0):
        self.group_size = group_size
        self.num_bits = num_bits
        self.q_int8 = q_int8
        self.num_groups = num_groups

    def quantize(self, inputs, qkv=True, count=1, parallel_dim=0):
        if not self.q_int8 or not qkv:
            inputs = torch.nn.Parameter(inputs, requires_grad=False)
            inputs.scale = torch.empty(1)
            return inputs
        q_range = 2 ** self.num_bits
        num_groups = self.num_groups if self.num_groups > 0 else inputs.shape[0] // self.group_size
        inputs = inputs.to(torch.float32)
        input_flat = inputs.reshape(num_groups, -1).contiguous()
        input_min = torch.min(input_flat, dim=1, keepdim=True)[0].float()
        input_max = torch.max(input_flat, dim=1, keepdim=True)[0].float()
        scale = torch.max(input_min.abs(), input_max.abs()) * 2.0 / q_range
        input_flat = (input_flat / scale).round().clamp(-q_range // 2, q_range // 2 - 1)
        out = torch.nn.Parameter(input_flat.reshape(inputs.shape).to(torch.int8).contiguous(), requires_grad=False)
        out.scale = scale.squeeze(1)
        return out


def replace_transformer_layer(orig_layer_impl, model, checkpoint_dict, config, model_config):
    """AutoTP / injection-policy entry point.  `config`: DeepSpeedInferenceConfig."""
    mp_size = config.tensor_parallel.tp_size
    mp_group = config.tensor_parallel.tp_group

    def replace_wo_policy(module, all_reduce_linears, prefix="", state_dict=None):
        _autotp = AutoTP(module, all_reduce_linears, prefix, state_dict, None, orig_layer_impl)
        _autotp.set_tensor_parallel_config(mp_size, mp_group)
        _autotp.update_linear_policies()
        return _autotp._replace_module(module)

    def replace_fn(child, _policy, layer_id=0, prefix="", state_dict=None):
        if isinstance(_policy, (tuple, list, set)) and not isinstance(_policy, str):
            return replace_wo_policy(child, list(_policy), prefix=prefix, state_dict=state_dict)
        return replace_wo_policy(child, [_policy], prefix=prefix, state_dict=state_dict)

    policy = {}
    if config.injection_policy_tuple is not None:
        for p in (config.injection_policy_tuple if isinstance(config.injection_policy_tuple[0], (tuple, list)) else [config.injection_policy_tuple]):
            policy[p[0]] = (replace_fn, p[1])
    replaced = replace_module(model=model, orig_class=orig_layer_impl, replace_fn=replace_fn, _replace_policy=config.injection_policy_tuple)
    # lm head (vocab projection) -> sharded over input features with all-reduce
    _autotp = AutoTP(model, [], "", None, None, orig_layer_impl)
    _autotp.set_tensor_parallel_config(mp_size, mp_group)
    _autotp._replace_last_linear_module(model)
    return replaced


def _client_returns_tuple(layer):
    """Does the client layer's forward return a tuple or the bare hidden states?  HF changed this per architecture over time and
    the surrounding model code indexes (or does not index) the result accordingly, so the replacement must return the same kind.
    Decided from the last `return` statement of the client's forward; None when the source is unavailable."""
    import inspect
    try:
        src = inspect.getsource(type(layer).forward)
    except (OSError, TypeError):
        return None
    returns = [ln.strip() for ln in src.splitlines() if ln.strip().startswith("return ")]
    if not returns:
        return None
    expr, depth = returns[-1][len("return "):].split("#")[0], 0
    for ch in expr:
        depth += ch in "([{"
        depth -= ch in ")]}"
        if ch == "," and depth == 0:
            return True
    return expr.strip().startswith("(") and expr.strip().endswith(",)")


def replace_with_kernel_policies(model, config, model_config):
    """Kernel injection, layer by layer: every layer whose class a policy of `replace_policies` recognises is rebuilt as a
    DeepSpeedTransformerInference block through that policy's container (reference replace_module.py:182-390 replace_with_policy).
    Returns the number of converted layers."""
    from .policy import TransformerPolicy
    from .replace_policy import replace_policies
    from .utils import policy_to_ds_container
    mp_size, mp_group = config.tensor_parallel.tp_size, config.tensor_parallel.tp_group
    mp_replace = ReplaceWithTensorSlicing(mp_group=mp_group, mp_size=mp_size)
    TransformerPolicy.hf_model_config = model_config
    table = {}
    for policy_cls in replace_policies:
        try:
            probe = policy_cls(None)  # resolves the client layer class (imports transformers lazily)
        except Exception:
            continue
        classes = policy_cls._orig_layer_class if policy_cls._orig_layer_class is not None else getattr(probe, "_orig_layer_class", None)
        for c in (classes if isinstance(classes, (list, tuple)) else [classes]):
            if c is not None:
                table.setdefault(c, policy_cls)
    converted = [0]

    def build(child, policy_cls, layer_id):
        policy = policy_cls(child, inference=True)
        container = policy_to_ds_container(policy=policy, config=config, model_config=model_config, layer_id=layer_id, child=child)
        container.set_tensor_parallel_config(mp_size, mp_group)
        container.create_ds_model_config()
        container.create_module()
        container.initialize_tensors()
        if getattr(container, "is_meta", False):
            # the client lives on the meta device: the block is filled later from a checkpoint (load_checkpoint.py); start from
            # zero biases and identity norms so tensors the checkpoint does not have (GPT-Neo / GPT-J attention biases) are neutral
            with torch.no_grad():
                for n, prm in container.module.named_parameters():
                    prm.fill_(1.0) if n.endswith(("norm_w", "attn_nw")) else prm.zero_()
        container.transpose()
        container.apply_tensor_parallelism(mp_replace)
        container.copy_data_to_new_module()
        container.module.config.layer_id = layer_id
        as_tuple = _client_returns_tuple(child)
        if as_tuple is not None:
            container.module.config.return_tuple = as_tuple
            container.module.config.return_single_tuple = False
            container.module.tensor_output_only = not as_tuple
        return container.module

    def walk(module):
        for name, child in module.named_children():
            policy_cls = table.get(child.__class__)
            if policy_cls is not None:
                setattr(module, name, build(child, policy_cls, converted[0]))
                converted[0] += 1
            else:
                walk(child)

    walk(model)
    return converted[0]


def replace_module(model, orig_class, replace_fn, _replace_policy, checkpoint=None):
    policy = {}
    if orig_class is not None:
        policy.update({orig_class: (replace_fn, _replace_policy)})
    else:
        for plcy in _replace_policy:
            policy.update({plcy[0]: (replace_fn, plcy[1])})
    replaced_module, _ = _replace_module(model, policy)
    return replaced_module


def _replace_module(model, policies, prefix="", layer_id=0, level_id=0, state_dict=None):
    for name, child in model.named_children():
        if child.__class__ in policies:
            replaced_module = policies[child.__class__][0](child, policies[child.__class__][-1], layer_id, prefix=prefix + name,
                                                           state_dict=state_dict)
            setattr(model, name, replaced_module)
            layer_id += 1
        else:
            _, layer_id = _replace_module(child, policies, prefix if level_id == 0 and not any(c.isdigit() for c in name) else prefix + name + ".",
                                          layer_id=layer_id, level_id=level_id + 1, state_dict=state_dict)
    return model, layer_id


def revert_transformer_layer(orig_layer_impl, model, config, preln=False):
    raise NotImplementedError("reverting injected layers is only meaningful for the reference's BERT training kernel injection")


def generic_injection(module, dtype=None, enable_cuda_graph=True):
    """Diffusers pipelines (reference replace_module.py:85-180): sub-modules matched by a generic policy (UNet, VAE) are wrapped
    in their CUDA-graph replay wrappers; a CLIP text encoder is wrapped in DSClipEncoder.  Returns the module (pipeline) itself."""
    from .replace_policy import generic_policies
    policies = [p() for p in generic_policies]
    for name in ("unet", "vae", "text_encoder"):
        sub = getattr(module, name, None)
        if sub is None:
            continue
        for policy in policies:
            if policy.match(sub):
                setattr(module, name, policy.apply(sub, enable_cuda_graph=enable_cuda_graph))
                break
        else:
            if name == "text_encoder" and type(sub).__name__.startswith("CLIPText"):
                from ..model_implementations.transformers.clip_encoder import DSClipEncoder
                setattr(module, name, DSClipEncoder(sub, enable_cuda_graph=enable_cuda_graph))
    return module
