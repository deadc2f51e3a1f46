"""CUDA-graph replay around a module's forward: one graph per distinct input signature (shapes / dtypes / non-tensor
arguments), static input buffers refilled by copy, outputs returned as the graph's static tensors.

The diffusion wrappers (UNet, VAE, CLIP text encoder) call the same shapes hundreds of times per image, so their forward is
launch-bound; a captured graph removes the per-kernel launch cost.  Without a CUDA device the wrapper runs eagerly.
(reference: deepspeed/model_implementations/features/cuda_graph.py + diffusers/unet.py:_create_cuda_graph / _graph_replay)"""
import torch


def _signature(args, kwargs):
    def one(v):
        if torch.is_tensor(v):
            return ("T", tuple(v.shape), v.dtype, v.device.type)
        if isinstance(v, (list, tuple)):
            return tuple(one(x) for x in v)
        if isinstance(v, dict):
            return tuple(sorted((k, one(x)) for k, x in v.items()))
        return ("V", repr(v))

    return (tuple(one(a) for a in args), tuple(sorted((k, one(v)) for k, v in kwargs.items())))


def _clone_static(v):
    if torch.is_tensor(v):
        return v.clone()
    if isinstance(v, (list, tuple)):
        return type(v)(_clone_static(x) for x in v)
    if isinstance(v, dict):
        return {k: _clone_static(x) for k, x in v.items()}
    return v


def _refill(static, new):
    if torch.is_tensor(static):
        static.copy_(new)
    elif isinstance(static, (list, tuple)):
        for s, n in zip(static, new):
            _refill(s, n)
    elif isinstance(static, dict):
        for k in static:
            _refill(static[k], new[k])


class CUDAGraph(object):
    """Mixin: subclasses implement `_forward(*args, **kwargs)`; `forward` replays a captured graph when enabled."""

    WARMUP_ITERS = 3

    def __init__(self, enable_cuda_graph=True):
        self.enable_cuda_graph = bool(enable_cuda_graph) and torch.cuda.is_available()
        self._graphs = {}
        self.cuda_graph_created = False

    def _graphed_call(self, fwd, *args, **kwargs):
        if not self.enable_cuda_graph or not any(torch.is_tensor(a) and a.is_cuda for a in list(args) + list(kwargs.values())):
            return fwd(*args, **kwargs)
        key = (fwd.__name__, _signature(args, kwargs))
        entry = self._graphs.get(key)
        if entry is None:
            entry = self._capture(fwd, args, kwargs)
            self._graphs[key] = entry
            self.cuda_graph_created = True
        graph, s_args, s_kwargs, s_out = entry
        _refill(s_args, args)
        _refill(s_kwargs, kwargs)
        graph.replay()
        return s_out

    def _capture(self, fwd, args, kwargs):
        s_args, s_kwargs = _clone_static(args), _clone_static(kwargs)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(self.WARMUP_ITERS):  # workspaces, lazy handles and autotuners settle before capture
                fwd(*s_args, **s_kwargs)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph), torch.no_grad():
            s_out = fwd(*s_args, **s_kwargs)
        return graph, s_args, s_kwargs, s_out
