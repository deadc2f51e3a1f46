"""Largest Hessian eigenvalue per transformer block by power iteration (Hessian-vector products through double
backward).  MoQ uses the normalised values to stretch the quantisation period of high-curvature blocks.

    v <- random unit vector over the block's parameters
    repeat:  Hv = d(g . v)/dw ;  lambda = <Hv, v> ;  v = Hv / ||Hv||     until |d lambda / lambda| < tol

The caller must have run backward(create_graph=True) so `param.grad` carries a graph.  Returns
{id(param): (eigenvalue scaled to [0, 1], block index)}.  (API parity: deepspeed/runtime/eigenvalue.py)"""
import torch

from ..utils.logging import log_dist


def _dot(xs, ys):
    return sum((x * y).sum() for x, y in zip(xs, ys))


def _finite(t):
    return torch.nan_to_num(t, nan=0.0, posinf=0.0, neginf=0.0)


class Eigenvalue(object):

    def __init__(self, verbose=False, max_iter=100, tol=1e-2, stability=0, gas_boundary_resolution=1, layer_name="", layer_num=0):
        assert len(layer_name) > 0 and layer_num > 0
        self.verbose = verbose
        self.max_iter = max_iter
        self.tol = tol
        self.stability = stability
        self.gas_boundary_resolution = gas_boundary_resolution
        self.layer_name = layer_name
        self.layer_num = layer_num
        log_dist(f"enabled eigenvalue with verbose={verbose}, max_iter={max_iter}, tol={tol}, stability={stability}, "
                 f"gas_boundary_resolution={gas_boundary_resolution}, layer_name={layer_name}, layer_num={layer_num}", ranks=[0])

    # kept public: user code calls these helpers
    def nan_to_num(self, x):
        return _finite(x)

    def inner_product(self, xs, ys):
        return _dot(xs, ys)

    def normalize(self, v):
        norm = _dot(v, v)**0.5 + self.stability
        return [_finite(x / norm) for x in v]

    def get_layers(self, module):
        m = module
        for name in self.layer_name.split("."):
            assert hasattr(m, name), "layer_name configuration is invalid."
            m = getattr(m, name)
        return m

    def _block_eigenvalue(self, block, device, scale):
        """Power iteration on one block; None when the block has no double-differentiable gradients."""
        params = [p for p in block.parameters() if p.grad is not None and p.grad.grad_fn is not None]
        if not params:
            return None, 0
        grads = [p.grad for p in params]
        # the probe vector must not disturb the training RNG stream
        rng_state = torch.random.get_rng_state()
        v = [torch.randn(p.size(), device=device) if device is not None else torch.randn(p.size()) for p in params]
        torch.random.set_rng_state(rng_state)
        v = [x.to(p.dtype) for x, p in zip(self.normalize(v), params)]
        current, previous, it = 1.0, 0.0, 0
        while it < self.max_iter and abs(current) > 0 and abs((current - previous) / current) >= self.tol:
            previous = current
            Hv = torch.autograd.grad(grads, params, grad_outputs=v, only_inputs=True, retain_graph=True)
            Hv = [_finite(h).float() for h in Hv]
            current = _dot(Hv, [x.float() for x in v]).item()
            v = [(x / scale).to(p.dtype) for x, p in zip(self.normalize(Hv), params)]
            it += 1
        return current * scale, it

    def compute_eigenvalue(self, module, device=None, scale=1.0):
        layers = self.get_layers(module)
        values, keys = [], []
        for b in range(self.layer_num):
            block = layers[b]
            value, iters = self._block_eigenvalue(block, device, scale)
            if value is None:
                # e.g. fused transformer kernels without a double-backward
                log_dist("The model does NOT support eigenvalue computation.", ranks=[0], level=30)
                return []
            values.append(value)
            keys.append([id(p) for p in block.parameters()])
            if self.verbose:
                log_dist(f"block: {b}, power iteration: {iters}, eigenvalue: {value}", ranks=[0])
        values = self.post_process(values)
        if self.verbose:
            log_dist(f"post processed block_eigenvalue: {values}", ranks=[0])
        return {pid: (value, b) for b, (ids, value) in enumerate(zip(keys, values)) for pid in ids}

    def post_process(self, value_list):
        """|lambda| / max|lambda|; blocks that produced 0 (no valid fp16 estimate) get the most conservative value 1.0."""
        top = max(abs(v) for v in value_list)
        return [abs(v) / top if v != 0.0 else 1.0 for v in value_list]
