"""Selective activation recomputation: what a checkpointed region keeps instead of recomputing.

Under per-layer activation checkpointing the layer's forward runs again in backward.  On a 180 GB GPU most of that second
forward is avoidable: the outputs that are expensive to recompute per byte kept (the attention core, the GEMM outputs)
can be held across the region, and only the cheap memory-bound ops (norms, RoPE, gated activation) are re-run.  A region
is a pair of contexts handed to `torch.utils.checkpoint(..., context_fn=...)`:

    record (first forward) : ops called with a `tag` the policy keeps put their output into the region's store
    replay (recomputation) : the same ops pop their output from the store and skip their kernel; ops tagged
                             `unused_in_recompute` (an output no backward formula reads, e.g. the layer's last GEMM)
                             skip their kernel and return an uninitialised tensor

Policies (`LlamaConfig.checkpoint_policy`):
    "full"       recompute everything (reference behaviour, activation_checkpointing/checkpointing.py)
    "attention"  keep the attention core's (out, lse)
    "gemm"       keep attention + every GEMM output of the layer (qkv, o, gate_up); only norms / rope / swiglu re-run
    "none"       no checkpointing: every activation autograd saves stays resident (nothing is recomputed)
    "auto"       "none" if all saved activations fit in 60 % of the free device memory at first use, else "gemm" if the
                 kept bytes fit in half of it, else "attention"
The last GEMM of a layer is never recomputed under any policy but "full" (its output is not needed by backward).
"""
import torch

ATTN, QKV, O_PROJ, GATE_UP = "attn", "qkv", "o", "gate_up"
POLICIES = {
    "full": frozenset(),
    "attention": frozenset({ATTN}),
    "gemm": frozenset({ATTN, QKV, O_PROJ, GATE_UP}),
}

_REGION = None  # the region executing on this thread


class Region(object):

    def __init__(self, mode, store, keep, skip_unused):
        self.mode, self.store, self.keep, self.skip_unused = mode, store, keep, skip_unused

    def __enter__(self):
        global _REGION
        self._prev = _REGION
        _REGION = self
        return self

    def __exit__(self, *exc):
        global _REGION
        _REGION = self._prev
        return False


def current():
    return _REGION


def contexts(policy="attention"):
    """-> (forward context, recompute context) sharing one store, for torch.utils.checkpoint(context_fn=...)."""
    keep = POLICIES[policy]
    store = {}
    skip = policy != "full"
    return Region("record", store, keep, skip), Region("replay", store, keep, skip)


def replayed(tag):
    """The kept output for `tag` during recomputation, else None."""
    r = _REGION
    if r is not None and r.mode == "replay" and tag is not None:
        q = r.store.get(tag)
        if q:
            return q.pop(0)
    return None


def record(tag, value):
    r = _REGION
    if r is not None and r.mode == "record" and tag in r.keep:
        r.store.setdefault(tag, []).append(value)


def output_unused():
    """True while recomputing under a policy that may drop outputs no backward formula reads."""
    r = _REGION
    return r is not None and r.mode == "replay" and r.skip_unused


def resolve_policy(policy, kept_bytes_if_gemm, bytes_if_none=None):
    if policy != "auto":
        assert policy in POLICIES or policy == "none", f"unknown checkpoint_policy {policy}"
        return policy
    if torch.cuda.is_available():
        free, _ = torch.cuda.mem_get_info()
        free += torch.cuda.memory_reserved() - torch.cuda.memory_allocated()  # cached blocks are reusable
        if bytes_if_none is not None and bytes_if_none <= 0.6 * free:
            return "none"
        return "gemm" if kept_bytes_if_gemm <= free // 2 else "attention"
    return "attention"
