"""FLOPs / MACs / latency / parameter profiler.

Design (B200-native, differs from the reference's functional-patching approach at
deepspeed/profiling/flops_profiler/profiler.py:30-1260):
  * torch ops are counted at the ATen level by a TorchDispatchMode (so functional calls, tensor methods and modules
    are all seen once, including inside autograd Functions);
  * this framework's own CUDA kernels do not go through the dispatcher -- they report themselves through
    `record_flops()` (one `if` when no profiler is active), see ops/gemm.py and ops/flash_attention.py;
  * module attribution uses a forward pre/post hook pair that maintains a module stack: every counted op is charged to
    all modules on the stack, so parents aggregate their children with no post-pass;
  * latency per module is measured with CUDA events when the module runs on a GPU (one sync at stop_profile, not per
    module) and with perf_counter on CPU.
The public surface (FlopsProfiler, get_model_profile, the *_to_string helpers, print_model_profile layout) follows the
reference so existing scripts keep working.
"""
import sys
import time
from collections import OrderedDict
from functools import partial
from typing import List, Optional

import numpy as np
import torch
from torch.utils._python_dispatch import TorchDispatchMode

_ACTIVE: List["FlopsProfiler"] = []
DEFAULT_PRECISION = 2


def record_flops(flops, macs=None, name="custom"):
    """Called by this framework's native-kernel wrappers (GEMM, flash attention, paged attention ...)."""
    if not _ACTIVE:
        return
    macs = flops // 2 if macs is None else macs
    for p in _ACTIVE:
        p._charge(name, int(flops), int(macs))


def _prod(xs):
    r = 1
    for x in xs:
        r *= int(x)
    return r


def _mm_flops(args, out):
    a, b = args[0], args[1]
    m, k = a.shape[-2], a.shape[-1]
    n = b.shape[-1]
    batch = _prod(out.shape[:-2]) if out.dim() > 2 else 1
    macs = batch * m * n * k
    return 2 * macs, macs


def _addmm_flops(args, out):
    a, b = args[1], args[2]
    macs = a.shape[0] * a.shape[1] * b.shape[1]
    return 2 * macs + out.numel(), macs


def _baddbmm_flops(args, out):
    a, b = args[1], args[2]
    macs = a.shape[0] * a.shape[1] * a.shape[2] * b.shape[2]
    return 2 * macs + out.numel(), macs


def _conv_flops(args, out):
    x, w = args[0], args[1]
    groups = args[8] if len(args) > 8 else 1
    transposed = args[6] if len(args) > 6 else False
    kernel = _prod(w.shape[2:])
    cin_per_group = w.shape[0] // groups if transposed else w.shape[1]
    spatial = _prod(x.shape[2:]) if transposed else _prod(out.shape[2:])
    cout = out.shape[1]
    macs = x.shape[0] * spatial * kernel * cin_per_group * (cout if not transposed else w.shape[1] * groups)
    bias = out.numel() if (len(args) > 2 and args[2] is not None) else 0
    return 2 * macs + bias, macs


def _sdpa_flops(args, out):
    q, k, v = args[0], args[1], args[2]
    b = _prod(q.shape[:-2])
    sq, d = q.shape[-2], q.shape[-1]
    sk, dv = k.shape[-2], v.shape[-1]
    macs = b * sq * sk * (d + dv)
    return 2 * macs, macs


def _ew(mult):

    def f(args, out):
        o = out[0] if isinstance(out, (tuple, list)) else out
        return mult * (o.numel() if torch.is_tensor(o) else 0), 0

    return f


def _norm_flops(args, out):
    return 5 * args[0].numel(), 0


def _softmax_flops(args, out):
    return 3 * args[0].numel(), 0  # exp + sum + div, as upstream counts it coarse-grained


def _embedding_flops(args, out):
    return 0, 0


aten = torch.ops.aten
_OP_TABLE = {
    aten.mm.default: _mm_flops, aten.bmm.default: _mm_flops, aten.matmul.default: _mm_flops, aten.addmm.default: _addmm_flops,
    aten.baddbmm.default: _baddbmm_flops, aten.convolution.default: _conv_flops, aten._convolution.default: _conv_flops,
    aten._scaled_dot_product_flash_attention.default: _sdpa_flops, aten._scaled_dot_product_efficient_attention.default: _sdpa_flops,
    aten._scaled_dot_product_cudnn_attention.default: _sdpa_flops, aten._scaled_dot_product_flash_attention_for_cpu.default: _sdpa_flops,
    aten._scaled_dot_product_attention_math.default: _sdpa_flops,
    aten.native_layer_norm.default: _norm_flops, aten.native_batch_norm.default: _norm_flops, aten.native_group_norm.default: _norm_flops,
    aten._native_batch_norm_legit.default: _norm_flops, aten._native_batch_norm_legit_no_training.default: _norm_flops,
    aten._softmax.default: _softmax_flops, aten._log_softmax.default: _softmax_flops, aten.embedding.default: _embedding_flops,
}
for _name, _mult in (("add", 1), ("sub", 1), ("mul", 1), ("div", 1), ("relu", 1), ("gelu", 8), ("silu", 4), ("sigmoid", 4), ("tanh", 4),
                     ("exp", 1), ("pow", 1), ("sqrt", 1), ("rsqrt", 1), ("mean", 1), ("sum", 1), ("leaky_relu", 1), ("elu", 1),
                     ("hardtanh", 1), ("upsample_nearest2d", 1), ("upsample_bilinear2d", 4), ("avg_pool2d", 1), ("max_pool2d_with_indices", 1)):
    _pkt = getattr(aten, _name, None)
    if _pkt is None:
        continue
    for _ov in _pkt.overloads():
        _OP_TABLE.setdefault(getattr(_pkt, _ov), _ew(_mult))


class _CountMode(TorchDispatchMode):

    def __init__(self, prof):
        super().__init__()
        self.prof = prof

    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        out = func(*args, **(kwargs or {}))
        fn = _OP_TABLE.get(func)
        if fn is not None:
            try:
                o = out[0] if isinstance(out, (tuple, list)) else out
                flops, macs = fn(args, o)
                self.prof._charge(func.__name__ if hasattr(func, "__name__") else str(func), int(flops), int(macs))
            except Exception:
                pass
        return out


class FlopsProfiler(object):
    """Measures the latency, number of estimated floating-point operations and parameters of each module in a model.

    with engine: enabled by the `flops_profiler` config block, profile printed at `profile_step`.
    standalone : prof = FlopsProfiler(model); prof.start_profile(); model(x); prof.stop_profile();
                 prof.print_model_profile(); prof.end_profile()
    """

    def __init__(self, model, ds_engine=None, recompute_fwd_factor=0.0):
        self.model = model
        self.ds_engine = ds_engine
        self.recompute_fwd_factor = recompute_fwd_factor
        self.started = False
        self.func_patched = False
        self._mode = None
        self._stack = []
        self._handles = []
        self._pending_events = []

    # -- accounting --------------------------------------------------------------------------------------------
    def _charge(self, name, flops, macs):
        for m in self._stack:
            m.__flops__ += flops
            m.__macs__ += macs
        if self._stack:
            ops = self._stack[-1].__ops__
            e = ops.get(name)
            if e is None:
                ops[name] = [1, flops, macs]
            else:
                e[0] += 1
                e[1] += flops
                e[2] += macs
        else:
            self._orphan_flops += flops
            self._orphan_macs += macs

    def start_profile(self, ignore_list=None):
        self.reset_profile()
        self._orphan_flops = self._orphan_macs = 0

        def pre_hook(module, inputs):
            self._stack.append(module)
            if any(torch.is_tensor(t) and t.is_cuda for t in inputs):
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                module.__start_ev__ = ev
            else:
                module.__start_time__ = time.perf_counter()

        def post_hook(module, inputs, output):
            if self._stack and self._stack[-1] is module:
                self._stack.pop()
            ev = getattr(module, "__start_ev__", None)
            if ev is not None:
                end = torch.cuda.Event(enable_timing=True)
                end.record()
                self._pending_events.append((module, ev, end))
                module.__start_ev__ = None
            elif getattr(module, "__start_time__", None) is not None:
                module.__duration__ += time.perf_counter() - module.__start_time__
                module.__start_time__ = None

        def register(module):
            if ignore_list and type(module) in ignore_list:
                return
            self._handles.append(module.register_forward_pre_hook(pre_hook))
            self._handles.append(module.register_forward_hook(post_hook))

        self.model.apply(register)
        self._mode = _CountMode(self)
        self._mode.__enter__()
        _ACTIVE.append(self)
        self.started = True
        self.func_patched = True

    def stop_profile(self):
        if self.started and self.func_patched:
            if self in _ACTIVE:
                _ACTIVE.remove(self)
            if self._mode is not None:
                self._mode.__exit__(None, None, None)
                self._mode = None
            self.func_patched = False
        if self._pending_events:
            torch.cuda.synchronize()
            for module, s, e in self._pending_events:
                module.__duration__ += s.elapsed_time(e) / 1e3
            self._pending_events = []
        for h in self._handles:
            h.remove()
        self._handles = []
        self._stack = []

    def reset_profile(self):

        def add_or_reset_attrs(module):
            module.__flops__ = 0
            module.__macs__ = 0
            module.__ops__ = {}
            module.__params__ = sum(getattr(p, "ds_numel", p.numel()) for p in module.parameters(recurse=False))
            module.__start_time__ = None
            module.__start_ev__ = None
            module.__duration__ = 0

        self.model.apply(add_or_reset_attrs)

    def end_profile(self):
        if not self.started:
            return
        self.stop_profile()
        self.started = False

        def remove_profile_attrs(module):
            for a in ("__flops__", "__macs__", "__ops__", "__params__", "__start_time__", "__start_ev__", "__duration__"):
                if hasattr(module, a):
                    delattr(module, a)

        self.model.apply(remove_profile_attrs)

    # -- queries ------------------------------------------------------------------------------------------------
    def get_total_flops(self, as_string=False):
        total_flops = self.model.__flops__ + getattr(self, "_orphan_flops", 0)
        return number_to_string(total_flops) if as_string else total_flops

    def get_total_macs(self, as_string=False):
        total_macs = self.model.__macs__ + getattr(self, "_orphan_macs", 0)
        return macs_to_string(total_macs) if as_string else total_macs

    def get_total_duration(self, as_string=False):
        total_duration = self.model.__duration__
        return duration_to_string(total_duration) if as_string else total_duration

    def get_total_params(self, as_string=False):
        total_params = _subtree_params(self.model)
        return params_to_string(total_params) if as_string else total_params

    def is_expert_tensor_parallelism_enabled(self):
        for _, module in self.model.named_modules():
            if hasattr(module, "enable_expert_tensor_parallelism"):
                return module.enable_expert_tensor_parallelism
        return False

    def print_model_profile(self, profile_step=1, module_depth=-1, top_modules=1, detailed=True, output_file=None):
        if not self.started:
            return
        original_stdout = None
        f = None
        if output_file and output_file != "":
            import os
            dir_path = os.path.dirname(os.path.abspath(output_file))
            os.makedirs(dir_path, exist_ok=True)
            original_stdout = sys.stdout
            f = open(output_file, "w")
            sys.stdout = f
        try:
            total_flops = self.get_total_flops()
            total_macs = self.get_total_macs()
            total_duration = self.get_total_duration()
            total_params = self.get_total_params()
            self.flops, self.macs, self.params = total_flops, total_macs, total_params
            print("\n-------------------------- DeepSpeed Flops Profiler --------------------------")
            print(f"Profile Summary at step {profile_step}:")
            print("Notations:\ndata parallel size (dp_size), model parallel size(mp_size),\nnumber of parameters (params), number of multiply-accumulate operations(MACs),\n"
                  "number of floating-point operations (flops), floating-point operations per second (FLOPS),\nfwd latency (forward propagation latency), "
                  "bwd latency (backward propagation latency),\nstep (weights update latency), iter latency (sum of fwd, bwd and step latency)\n")
            line_fmt = "{:<70}  {:<8}"
            if self.ds_engine:
                print(line_fmt.format("world size: ", self.ds_engine.world_size))
                print(line_fmt.format("data parallel size: ", self.ds_engine.dp_world_size))
                print(line_fmt.format("model parallel size: ", self.ds_engine.mp_world_size))
                print(line_fmt.format("batch size per GPU: ", self.ds_engine.train_micro_batch_size_per_gpu()))
            print(line_fmt.format("params per GPU: ", params_to_string(total_params)))
            print(line_fmt.format("fwd MACs per GPU: ", macs_to_string(total_macs)))
            print(line_fmt.format("fwd flops per GPU: ", number_to_string(total_flops)))
            if self.ds_engine:
                print(line_fmt.format("fwd flops of model = fwd flops per GPU * mp_size: ", number_to_string(total_flops * (self.ds_engine.mp_world_size or 1))))
            fwd_latency = self.get_total_duration()
            if self.ds_engine and self.ds_engine.wall_clock_breakdown():
                from ...utils.timer import FORWARD_GLOBAL_TIMER
                fwd_latency = self.ds_engine.timers(FORWARD_GLOBAL_TIMER).elapsed(False) / 1000.0
            print(line_fmt.format("fwd latency: ", duration_to_string(fwd_latency)))
            print(line_fmt.format("fwd FLOPS per GPU = fwd flops per GPU / fwd latency: ", flops_to_string(total_flops / max(fwd_latency, 1e-12))))
            if self.ds_engine and self.ds_engine.wall_clock_breakdown():
                from ...utils.timer import BACKWARD_GLOBAL_TIMER, STEP_GLOBAL_TIMER
                bwd_factor = 2 + self.recompute_fwd_factor
                bwd_latency = self.ds_engine.timers(BACKWARD_GLOBAL_TIMER).elapsed(False) / 1000.0
                step_latency = self.ds_engine.timers(STEP_GLOBAL_TIMER).elapsed(False) / 1000.0
                print(line_fmt.format("bwd latency: ", duration_to_string(bwd_latency)))
                print(line_fmt.format(f"bwd FLOPS per GPU = {bwd_factor:g} * fwd flops per GPU / bwd latency: ",
                                      flops_to_string(bwd_factor * total_flops / max(bwd_latency, 1e-12))))
                print(line_fmt.format(f"fwd+bwd FLOPS per GPU = {bwd_factor + 1:g} * fwd flops per GPU / (fwd+bwd latency): ",
                                      flops_to_string((bwd_factor + 1) * total_flops / max(fwd_latency + bwd_latency, 1e-12))))
                print(line_fmt.format("step latency: ", duration_to_string(step_latency)))
                iter_latency = fwd_latency + bwd_latency + step_latency
                print(line_fmt.format("iter latency: ", duration_to_string(iter_latency)))
                print(line_fmt.format(f"FLOPS per GPU = {bwd_factor + 1:g} * fwd flops per GPU / iter latency: ",
                                      flops_to_string((bwd_factor + 1) * total_flops / max(iter_latency, 1e-12))))
                samples_per_iter = self.ds_engine.train_micro_batch_size_per_gpu() * self.ds_engine.world_size
                print(line_fmt.format("samples/second: ", round(samples_per_iter / max(iter_latency, 1e-12), DEFAULT_PRECISION)))

            def flops_repr(module):
                mod_params = _subtree_params(module)
                flops = module.__flops__
                macs = module.__macs__
                duration = module.__duration__
                items = ["{} = {:g}% Params".format(params_to_string(mod_params), round(100 * mod_params / total_params, DEFAULT_PRECISION) if total_params else 0),
                         "{} = {:g}% MACs".format(macs_to_string(macs), round(100 * macs / total_macs, DEFAULT_PRECISION) if total_macs else 0),
                         "{} = {:g}% latency".format(duration_to_string(duration), round(100 * duration / total_duration, DEFAULT_PRECISION) if total_duration else 0),
                         flops_to_string(round(flops / duration, DEFAULT_PRECISION) if duration else 0)]
                original_extra_repr = module.original_extra_repr()
                if original_extra_repr:
                    items.append(original_extra_repr)
                return ", ".join(items)

            def add_extra_repr(module):
                flops_extra_repr = flops_repr.__get__(module)
                if module.extra_repr != flops_extra_repr:
                    module.original_extra_repr = module.extra_repr
                    module.extra_repr = flops_extra_repr
                    assert module.extra_repr != module.original_extra_repr

            def del_extra_repr(module):
                if hasattr(module, "original_extra_repr"):
                    module.extra_repr = module.original_extra_repr
                    del module.original_extra_repr

            self.model.apply(add_extra_repr)
            print("\n----------------------------- Aggregated Profile per GPU -----------------------------")
            self.print_model_aggregated_profile(module_depth=module_depth, top_modules=top_modules)
            if detailed:
                print("\n------------------------------ Detailed Profile per GPU ------------------------------")
                print("Each module profile is listed after its name in the following order: \nparams, percentage of total params, MACs, percentage of total MACs, "
                      "fwd latency, percentage of total fwd latency, fwd FLOPS")
                print("\nNote: 1. A module can have torch.nn.module or torch.nn.functional to compute logits (e.g. CrossEntropyLoss). They are not counted as submodules, "
                      "thus not to be printed out. However they make up the difference between a parent's MACs (or latency) and the sum of its submodules'.\n"
                      "2. Number of floating-point operations is a theoretical estimation, thus FLOPS computed using that could be larger than the maximum system throughput.\n"
                      "3. The fwd latency listed in the top module's profile is directly captured at the module forward function in PyTorch, thus it's less than the fwd "
                      "latency shown above which is captured in DeepSpeed.\n")
                print(self.model)
            self.model.apply(del_extra_repr)
            print("------------------------------------------------------------------------------")
        finally:
            if f is not None:
                sys.stdout = original_stdout
                f.close()

    def print_model_aggregated_profile(self, module_depth=-1, top_modules=1):
        info = {}
        if not hasattr(self.model, "__flops__"):
            print("no __flops__ attribute in the model, call this function after start_profile and before end_profile")
            return

        def walk_module(module, curr_depth, info):
            if curr_depth not in info:
                info[curr_depth] = {}
            if module.__class__.__name__ not in info[curr_depth]:
                info[curr_depth][module.__class__.__name__] = [0, 0, 0]
            info[curr_depth][module.__class__.__name__][0] += _subtree_params(module)
            info[curr_depth][module.__class__.__name__][1] += module.__macs__
            info[curr_depth][module.__class__.__name__][2] += module.__duration__
            has_children = len(module._modules.items()) != 0
            if has_children:
                for child in module.children():
                    walk_module(child, curr_depth + 1, info)

        walk_module(self.model, 0, info)
        depth = module_depth
        if module_depth == -1:
            depth = len(info) - 1
        print(f"Top {top_modules} modules in terms of params, MACs or fwd latency at different model depths:")
        for d in range(depth):
            num_items = min(top_modules, len(info[d]))
            sort_params = {k: params_to_string(v[0]) for k, v in sorted(info[d].items(), key=lambda item: item[1][0], reverse=True)[:num_items]}
            sort_macs = {k: macs_to_string(v[1]) for k, v in sorted(info[d].items(), key=lambda item: item[1][1], reverse=True)[:num_items]}
            sort_time = {k: duration_to_string(v[2]) for k, v in sorted(info[d].items(), key=lambda item: item[1][2], reverse=True)[:num_items]}
            print(f"depth {d}:")
            print(f"    params      - {sort_params}")
            print(f"    MACs        - {sort_macs}")
            print(f"    fwd latency - {sort_time}")


def _subtree_params(module):
    return sum(getattr(m, "__params__", 0) for m in module.modules())


# ---- string helpers (same names and formats as the reference) ---------------------------------------------------------
def number_to_string(num, units=None, precision=DEFAULT_PRECISION):
    if units is None:
        if num >= 1e12:
            magnitude, units = 1e12, "T"
        elif num >= 1e9:
            magnitude, units = 1e9, "G"
        elif num >= 1e6:
            magnitude, units = 1e6, "M"
        elif num >= 1e3:
            magnitude, units = 1e3, "K"
        elif num >= 1 or num == 0:
            magnitude, units = 1, ""
        elif num >= 1e-3:
            magnitude, units = 1e-3, "m"
        else:
            magnitude, units = 1e-6, "u"
    else:
        magnitude = {"T": 1e12, "G": 1e9, "M": 1e6, "K": 1e3, "m": 1e-3, "u": 1e-6}.get(units, 1)
        if magnitude == 1:
            units = ""
    return f"{round(num / magnitude, precision):g} {units}"


def flops_to_string(flops, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(flops, units=units, precision=precision)}FLOPS"


def bytes_to_string(b, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(b, units=units, precision=precision)}B"


def macs_to_string(macs, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(macs, units=units, precision=precision)}MACs"


def params_to_string(params_num, units=None, precision=DEFAULT_PRECISION):
    units = units.replace("B", "G") if units else units
    return number_to_string(params_num, units=units, precision=precision).replace("G", "B").strip()


def duration_to_string(duration, units=None, precision=DEFAULT_PRECISION):
    return f"{number_to_string(duration, units=units, precision=precision)}s"


def get_model_profile(model, input_shape=None, args=[], kwargs={}, print_profile=True, detailed=True, module_depth=-1, top_modules=1, warm_up=1,
                      as_string=True, output_file=None, ignore_modules=None, mode="forward"):
    """Returns (flops, macs, params) of one forward (or generate) of `model`.  Parity: profiler.py:1160-1260."""
    assert isinstance(model, torch.nn.Module), "model must be a PyTorch module"
    prof = FlopsProfiler(model)
    model.eval()
    if input_shape is not None:
        assert type(input_shape) is tuple, "input_shape must be a tuple"
        assert len(input_shape) >= 1, "input_shape must have at least one element"
        try:
            input = torch.ones(()).new_empty((*input_shape, ), dtype=next(model.parameters()).dtype, device=next(model.parameters()).device)
        except StopIteration:
            input = torch.ones(()).new_empty((*input_shape, ))
        args = [input]
    assert (len(args) > 0) or (len(kwargs) > 0), "args and/or kwargs must be specified if input_shape is None"
    for _ in range(warm_up):
        with torch.no_grad():
            if mode == "forward":
                _ = model(*args, **kwargs)
            if mode == "generate":
                _ = model.generate(*args, **kwargs)
    prof.start_profile(ignore_list=ignore_modules)
    with torch.no_grad():
        if mode == "forward":
            _ = model(*args, **kwargs)
        elif mode == "generate":
            _ = model.generate(*args, **kwargs)
        else:
            raise NotImplementedError("mode must be 'forward' or 'generate'")
    prof.stop_profile()
    flops = prof.get_total_flops()
    macs = prof.get_total_macs()
    params = prof.get_total_params()
    if print_profile:
        prof.print_model_profile(profile_step=warm_up, module_depth=module_depth, top_modules=top_modules, detailed=detailed, output_file=output_file)
    prof.end_profile()
    if as_string:
        return number_to_string(flops), macs_to_string(macs), params_to_string(params)
    return flops, macs, params
