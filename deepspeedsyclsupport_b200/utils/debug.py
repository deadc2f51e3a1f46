"""Debug helpers: fully-qualified module / parameter names for log messages and rank-gated printing
(parity: deepspeed/utils/debug.py)."""
import fcntl
import os

module_names = {}
param_names = {}


def debug_clear_module_and_param_names():
    global module_names, param_names
    module_names = {}
    param_names = {}


def debug_extract_module_and_param_names(model):
    global module_names, param_names
    module_names = {module: name for name, module in model.named_modules()}
    param_names = {param: name for name, param in model.named_parameters()}


def debug_module2name(module):
    return module_names.get(module, "unknown")


def debug_module2name_id(module):
    return f"name={debug_module2name(module)} id={getattr(module, 'id', None)}"


def debug_module2name_class(module):
    return f"name={debug_module2name(module)} {module.__class__.__name__}"


def debug_param2name(param):
    return param_names.get(param, getattr(param, "ds_name", "unknown"))


def debug_param2name_id(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)}"


def debug_param2name_id_shape(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)} shape={tuple(param.data.shape)}"


def debug_param2name_id_shape_device(param):
    return f"{debug_param2name_id_shape(param)} device={param.device}"


def debug_param2name_id_numel(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', None)} numel={param.numel()}"


def debug_param2name_id_shape_status(param):
    return f"{debug_param2name_id_shape(param)} status={getattr(param, 'ds_status', None)}"


def printflock(*msgs):
    """print under an exclusive file lock so multi-process output lines do not interleave"""
    with open(__file__, "r") as fh:
        fcntl.flock(fh, fcntl.LOCK_EX)
        try:
            print(*msgs)
        finally:
            fcntl.flock(fh, fcntl.LOCK_UN)


fh = None


def log_rank_file(rank, *msgs):
    """one log file per rank in the current directory (log_rank_<r>.txt)"""
    global fh
    if fh is None:
        fh = open(f"log_rank_{rank}.txt", "w")
    for m in msgs:
        fh.write(f"{m}\n")
    fh.flush()


def print_backward_tensors(tensor):

    def _walk(fn):
        print(f"Backward tensors in {fn}")
        for funcs in fn.next_functions:
            if funcs[0]:
                try:
                    t = getattr(funcs[0], "variable")
                    print(funcs[0], f"Tensor - id: {id(t)}, shape: {t.shape}, data: {t}, grad: {t.grad}")
                except AttributeError:
                    _walk(funcs[0])

    if hasattr(tensor, "grad_fn"):
        _walk(tensor.grad_fn)
