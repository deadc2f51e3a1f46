"""DeepSpeedHybridEngine: one engine that trains (ZeRO) and generates (RLHF actor) with the SAME weights.

B200 design.  The reference swaps every transformer layer for an inference-kernel container that aliases the training
weights (deepspeed/runtime/hybrid_engine.py:32-446).  Here the training layers already run on this framework's kernels
(tcgen05 GEMM, fused norm / rope / attention), so generation keeps the module as is and the engine only manages what
differs between the two phases:
  * ZeRO-3: all parameter units are gathered ONCE before generate (pin_parameters) -- with 180 GB of HBM a 7B-70B actor
    fits gathered -- and released after, instead of being re-gathered for each of the hundreds of decode forwards;
    without pinning the usual per-layer fetch hooks serve every forward.
  * LoRA (DeepSpeed-Chat `LinearLayer_LoRA` convention: weight, lora_right_weight, lora_left_weight, lora_scaling) is
    folded into the base weight for the duration of generate and unfolded afterwards, so decode runs one GEMM per layer.
  * generate runs under no_grad / eval with the hooks' trace recording suspended; latency of gather / generate / train is
    tracked and printed by eval() like the reference.
  * inference_tp_size > 1 re-shards for generation; the all-gather of the per-rank prompts and the split of the outputs
    follow the reference semantics (each TP rank generates for the whole TP group's batch)."""
import time

import torch

from .. import comm as dist
from ..accelerator import get_accelerator
from ..utils.logging import log_dist, logger
from .engine import DeepSpeedEngine
from .zero.param_store import ZeroParamStatus


def _lora_modules(module):
    for m in module.modules():
        if hasattr(m, "lora_right_weight") and hasattr(m, "lora_left_weight") and hasattr(m, "weight"):
            yield m


class DeepSpeedHybridEngine(DeepSpeedEngine):
    inference_mp_group = None

    def __init__(self, args, model, **kwargs):
        super().__init__(args, model, **kwargs)
        if dist.is_initialized() and get_accelerator().is_available():
            # every rank must sample the same tokens when ranks share a prompt (TP generation)
            rng = get_accelerator().get_rng_state().to(get_accelerator().current_device_name())
            dist.broadcast(rng, 0)
            get_accelerator().set_rng_state(rng.cpu())
        self.Z3_enabled = self._config.zero_config.stage == 3
        self.gather_all_layers = self._config.hybrid_engine.pin_parameters
        self.is_lora_fused = False
        self._pinned = []
        self._t_start = None
        self._total_latency = 0.0
        self._iters = 0
        self._training_start_time = None
        self._generate_latency = 0.0
        self._training_latency = 0.0
        self._gather_latency = 0.0
        self._total_batch_size = None
        self.mp_group = None
        tp = self._config.hybrid_engine.inference_tp_size
        if tp > 1:
            self._create_inference_tp_groups(tp)
        if hasattr(self.module, "generate"):
            self._generate = self.module.generate
            self.module.generate = self.generate
        else:
            self._generate = None

    def _create_inference_tp_groups(self, tp):
        world = dist.get_world_size()
        assert world % tp == 0, f"inference_tp_size {tp} must divide the world size {world}"
        for start in range(0, world, tp):
            ranks = list(range(start, start + tp))
            group = dist.new_group(ranks)
            if dist.get_rank() in ranks:
                self.mp_group = group
        DeepSpeedHybridEngine.inference_mp_group = self.mp_group

    # ------------------------------------------------------------------ LoRA folding
    @torch.no_grad()
    def _apply_lora(self, sign):
        for m in _lora_modules(self.module):
            scaling = getattr(m, "lora_scaling", 1.0)
            delta = (m.lora_right_weight.float() @ m.lora_left_weight.float()).t() * scaling  # [out, in], DeepSpeed-Chat layout
            if delta.shape != m.weight.shape:
                delta = delta.t()
            m.weight.data.add_(delta.to(m.weight.dtype), alpha=sign)

    def fuse_lora_weight(self):
        if not self.is_lora_fused:
            self._apply_lora(+1.0)
            self.is_lora_fused = True

    def unfuse_lora_weight(self):
        if self.is_lora_fused:
            self._apply_lora(-1.0)
            self.is_lora_fused = False

    def unfuse_lora_weight_non_pinned(self):
        self.unfuse_lora_weight()

    # ------------------------------------------------------------------ ZeRO-3 pinning
    def _pin_all_units(self):
        opt = self.optimizer
        if not (self.Z3_enabled and opt is not None and getattr(opt, "shard_world", 1) > 1):
            return
        t0 = time.time()
        self._pinned = []
        for u in opt.units:
            if u.status != ZeroParamStatus.AVAILABLE:
                opt._fetch(u)
            if not u.persistent:
                u.persistent = True  # the forward hooks leave persistent units alone -> no release during decode
                self._pinned.append(u)
        if get_accelerator().is_available():
            get_accelerator().synchronize()
        self._gather_latency = time.time() - t0

    def _unpin_all_units(self):
        opt = self.optimizer
        for u in self._pinned:
            u.persistent = False
            opt.release_unit(u)
        self._pinned = []

    def retake_inference_cache(self):
        pass  # KV caches are per-generate tensors owned by the model; nothing is held between calls

    # ------------------------------------------------------------------ generate
    def generate(self, *inputs, **kwargs):
        assert self._generate is not None, "the wrapped module has no generate()"
        if self._total_batch_size is None and len(inputs) > 0 and torch.is_tensor(inputs[0]):
            self._total_batch_size = inputs[0].shape[0] * (dist.get_world_size() if dist.is_initialized() else 1)
        self._t0 = time.time()
        self._gather_latency = 0.0
        was_training = self.module.training
        self.module.eval()
        if self.gather_all_layers:
            self._pin_all_units()
        try:
            self.fuse_lora_weight()
            tp = self._config.hybrid_engine.inference_tp_size
            with torch.no_grad():
                if tp > 1 and self.mp_group is not None and len(inputs) > 0 and torch.is_tensor(inputs[0]):
                    # all TP ranks of a group generate for the concatenation of their prompts, then keep their own rows
                    local = inputs[0]
                    gathered = torch.empty(local.shape[0] * tp, *local.shape[1:], dtype=local.dtype, device=local.device)
                    dist.all_gather_into_tensor(gathered, local.contiguous(), group=self.mp_group)
                    out = self._generate(gathered, *inputs[1:], **kwargs)
                    r = dist.get_rank(group=self.mp_group)
                    out = out[local.shape[0] * r:local.shape[0] * (r + 1)]
                else:
                    out = self._generate(*inputs, **kwargs)
        finally:
            self.unfuse_lora_weight()
            if self.gather_all_layers:
                self._unpin_all_units()
            if was_training:
                self.module.train()
        if self._config.hybrid_engine.release_inference_cache and get_accelerator().is_available():
            get_accelerator().empty_cache()
        self._generate_latency = time.time() - self._t0 - self._gather_latency
        return out

    # ------------------------------------------------------------------ phase switches + stats
    def eval(self):
        if self._t_start is not None:
            latency = time.time() - self._t_start
            self._total_latency += latency
            self._iters += 1
            if not dist.is_initialized() or dist.get_rank() == 0:
                others = latency - (self._generate_latency + self._training_latency + self._gather_latency)
                pct = lambda x: x / max(latency, 1e-9) * 100
                print(f"|E2E latency={latency:.2f}s |Gather latency={self._gather_latency:.2f}s ({pct(self._gather_latency):.2f}%) "
                      f"|Generate time={self._generate_latency:.2f}s ({pct(self._generate_latency):.2f}%) "
                      f"|Training time={self._training_latency:.2f}s ({pct(self._training_latency):.2f}%) "
                      f"|Others={others:.2f} ({pct(others):.2f}%)|CurSamplesPerSec={(1 / max(latency, 1e-9) * (self._total_batch_size or 0)):.2f} "
                      f"|AvgSamplesPerSec={(1 / max(self._total_latency / self._iters, 1e-9) * (self._total_batch_size or 0)):.2f}")
            self._t_start = time.time()
        self._training_latency = 0.0
        super().eval()
        if self._t_start is None:
            self._t_start = time.time()

    def train(self, mode=True):
        if mode and self._t_start is None:
            self._t_start = time.time()
        self._training_start_time = time.time()
        super().train(mode)

    def step(self, lr_kwargs=None):
        super().step(lr_kwargs=lr_kwargs)
        if self._training_start_time is not None:
            self._training_latency += time.time() - self._training_start_time
            self._training_start_time = time.time()

    # ------------------------------------------------------------------ reference container surface
    # The reference builds one inference "container" per transformer layer that aliases the training weights into its
    # fused-kernel module (hybrid_engine.py:214-300).  Here the training layers already run on the fused kernels, so a
    # container is just a record of (layer, index, architecture): these methods keep the names for code that introspects
    # the hybrid engine, and `_inference_containers` lists the layers whose parameter units are pinned around generate().
    def populate_all_inference_policies(self):
        """Architectures generation knows about: the ragged-engine families by `model_type`, and every client layer class a
        layer policy of module_inject/replace_policy recognises (reference hybrid_engine.py:118-127)."""
        from ..inference.kernel_inject import SUPPORTED
        from ..module_inject.replace_policy import replace_policies
        self.inference_policies = {name: name for name in SUPPORTED}
        for policy_cls in replace_policies:
            try:
                policy_cls(None)
            except Exception:
                continue
            classes = policy_cls._orig_layer_class
            for c in (classes if isinstance(classes, (list, tuple)) else [classes]):
                if c is not None:
                    self.inference_policies[c] = policy_cls
        return self.inference_policies

    def new_inference_container(self, orig_layer, policy_cls, layer_id):
        policies = getattr(self, "inference_policies", {})
        return {"layer": orig_layer, "policy": policies.get(type(orig_layer), policy_cls), "layer_id": layer_id,
                "units": sorted({id(getattr(p, "ds_unit", None)) for p in orig_layer.parameters() if hasattr(p, "ds_unit")})}

    def create_inference_containers(self, module, layer_id=0):
        """Transformer blocks of `module` in execution order: the children of the longest ModuleList found."""
        lists = [m for m in module.modules() if isinstance(m, torch.nn.ModuleList) and len(m) > 0]
        blocks = max(lists, key=len) if lists else []
        policy = getattr(getattr(self.module, "config", None), "model_type", None)
        self._inference_containers = [self.new_inference_container(b, policy, layer_id + i) for i, b in enumerate(blocks)]
        return self._inference_containers

    def create_inference_module(self):
        """(Re)bind generation to this engine: containers are rebuilt and `module.generate` routes through `generate()`."""
        self.populate_all_inference_policies()
        self.create_inference_containers(self.module)
        if hasattr(self.module, "generate") and getattr(self.module.generate, "__self__", None) is not self:
            self._generate = self.module.generate
            self.module.generate = self.generate
        return self.module

    def convert_to_linear_transposed(self, model):
        """The reference transposes nn.Linear weights for its inference GEMMs; the tcgen05 GEMM takes [out, in] weights
        natively (K-major B operand), so nothing is rewritten."""
        return model
