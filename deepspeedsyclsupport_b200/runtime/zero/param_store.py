"""Flat sharded parameter store: the data layout under every ZeRO stage of this framework.

The reference partitions each parameter separately (`ceil(numel/world)` per rank, partition_parameters.py:1431-1548),
re-assembles a module's parameters with torch.cat/narrow copies on every fetch (:1064-1098, :618-657) and decides what
to fetch with a per-submodule hook + runtime trace (partitioned_param_coordinator.py:256-460).  Here parameters are
grouped into *units* (one per transformer block / leaf module); a unit is ONE contiguous low-precision buffer, padded
to `world * ALIGN`, and rank r owns the contiguous slice [r*S, (r+1)*S).  Consequences:

  * all-gather lands directly in the buffer the parameters are views of (zero copies, one collective per unit);
  * the gradient reduce-scatter reads one contiguous gradient buffer per unit and writes the owner's slice;
  * the optimizer sees one flat fp32 shard per rank -> a handful of fused-Adam launches per step;
  * with world == 1 every "gather" and "reduce" is the identity and parameters alias the shard.

The same store backs stage 0 (shard world = 1, gradients all-reduced), stages 1/2 (full buffers stay resident) and
stage 3 (full buffers are materialised per unit around its forward/backward, with trace-ordered prefetch on a side
stream), plus MiCS / hpZ style sub-group sharding (shard group != data-parallel group).
"""
import math
from enum import Enum

import torch

ALIGN = 128  # elements: shard boundaries are 256-byte aligned for 16-bit dtypes (TMA, 16-byte vectors) and whole fp8 quantisation blocks (ZeRO++)


class ZeroParamType(Enum):
    NORMAL = 1      # managed by the store
    PARTITIONED = 2
    REMOTE = 3


class ZeroParamStatus(Enum):
    AVAILABLE = 1       # full tensor materialised
    NOT_AVAILABLE = 2   # only the shard exists
    INFLIGHT = 3        # gather in progress


def _align_up(n, a):
    return (n + a - 1) // a * a


class ParamUnit:
    """A set of parameters that is gathered, released and gradient-reduced together."""

    def __init__(self, uid, name, params, group_ids, world, rank, dtype, device):
        self.uid = uid
        self.name = name
        # order by optimizer param-group so each unit contributes at most one run per group to the flat shard
        order = sorted(range(len(params)), key=lambda i: (group_ids[i], i))
        self.params = [params[i] for i in order]
        self.group_ids = [group_ids[i] for i in order]
        self.dtype = dtype
        self.device = device
        self.world, self.rank = world, rank
        self.offsets = []
        self.numels = []
        off = 0
        for p in self.params:
            n = int(p.ds_numel) if getattr(p, "ds_init_sharded", False) else p.numel()
            self.numels.append(n)
            self.offsets.append(off)
            off += _align_up(n, 8)  # 16-byte aligned starts for 16-bit dtypes
        self.numel = off
        self.padded = _align_up(max(off, 1), world * ALIGN)
        self.shard_numel = self.padded // world
        self.trainable = [p for p in self.params if p.requires_grad]
        # runtime state
        self.full = None         # [padded] model dtype, when materialised
        self.shard = None        # view into the rank's flat lp shard buffer
        self.shard_off = 0       # offset of this unit's shard inside the rank-flat buffers
        self.full_grad = None    # [padded] gradient buffer (comm dtype) while a backward is in flight
        self.status = ZeroParamStatus.NOT_AVAILABLE
        self.persistent = False
        self.gather_event = None
        self.ready = 0           # trainable params whose gradient is complete in this micro-step
        self.reduced = False
        self.modules = []        # modules whose forward needs this unit

    # ---- geometry helpers
    def param_slice(self, i):
        return self.offsets[i], self.numels[i]

    def shard_range(self):
        """[lo, hi) of the unit-flat index space owned by this rank."""
        return self.rank * self.shard_numel, (self.rank + 1) * self.shard_numel

    def group_runs(self):
        """Maximal runs [(group_id, lo, hi)] of the unit-flat space that belong to one optimizer param-group
        (alignment padding joins the preceding run)."""
        runs = []
        for i, g in enumerate(self.group_ids):
            lo = self.offsets[i]
            hi = self.offsets[i + 1] if i + 1 < len(self.params) else self.padded
            if runs and runs[-1][0] == g:
                runs[-1][2] = hi
            else:
                runs.append([g, lo, hi])
        if not runs:
            runs = [[0, 0, self.padded]]
        return [tuple(r) for r in runs]

    def shard_group_runs(self):
        """group_runs() clipped to this rank's shard, expressed in shard-local offsets."""
        lo, hi = self.shard_range()
        out = []
        for g, a, b in self.group_runs():
            a2, b2 = max(a, lo), min(b, hi)
            if a2 < b2:
                out.append((g, a2 - lo, b2 - lo))
        return out

    # ---- views
    def bind_params(self):
        """Point every parameter at its slice of `self.full`."""
        for i, p in enumerate(self.params):
            off, n = self.param_slice(i)
            p.data = self.full[off:off + n].view(p.ds_shape)
            p.ds_status = ZeroParamStatus.AVAILABLE
        self.status = ZeroParamStatus.AVAILABLE

    def unbind_params(self):
        for p in self.params:
            p.data = _empty_like_param(p)
            p.ds_status = ZeroParamStatus.NOT_AVAILABLE
        self.status = ZeroParamStatus.NOT_AVAILABLE

    def grad_views(self, buf):
        return [buf[self.offsets[i]:self.offsets[i] + self.numels[i]].view(p.ds_shape) for i, p in enumerate(self.params)]


_EMPTY_CACHE = {}


def _empty_like_param(p):
    key = (p.dtype, p.device)
    if key not in _EMPTY_CACHE:
        _EMPTY_CACHE[key] = torch.empty(0, dtype=p.dtype, device=p.device)
    return _EMPTY_CACHE[key]


def find_blocks(module):
    """Transformer-block discovery: direct children of every nn.ModuleList / nn.Sequential that own parameters and are
    not nested inside an already-selected block.  Returns [(qualified_name, module)]."""
    blocks, chosen = [], set()
    for name, m in module.named_modules():
        if any(id(a) in chosen for a in _ancestors_in(module, name)):
            continue
        if isinstance(m, (torch.nn.ModuleList, torch.nn.Sequential)) and m is not module:
            for cname, child in m.named_children():
                if sum(p.numel() for p in child.parameters()) > 0:
                    blocks.append((f"{name}.{cname}", child))
                    chosen.add(id(child))
    return blocks


def _ancestors_in(root, qualified_name):
    parts = qualified_name.split(".") if qualified_name else []
    out, cur = [], root
    for part in parts[:-1]:
        cur = cur._modules[part]
        out.append(cur)
    return out


def build_units(module, param_to_group, world, rank, dtype, device, max_unit_numel=None, param_filter=None):
    """Partition the parameters of `module` into units.  Shared (tied) parameters belong to the first unit that sees
    them.  Returns (units, param_id -> unit)."""
    seen = set()
    units, owner = [], {}

    def add_unit(name, plist, mods):
        plist = [p for p in plist if id(p) not in seen and (param_filter is None or param_filter(p))]
        if not plist:
            return None
        for p in plist:
            seen.add(id(p))
        u = ParamUnit(len(units), name, plist, [param_to_group.get(id(p), 0) for p in plist], world, rank, dtype, device)
        u.modules = list(mods)
        units.append(u)
        for p in plist:
            owner[id(p)] = u
        return u

    block_ids = set()
    blocks = find_blocks(module)
    for _, b in blocks:
        for sub in b.modules():
            block_ids.add(id(sub))
    # walk in registration order so units follow execution order for typical models
    block_by_id = {id(b): (n, b) for n, b in blocks}
    for name, m in module.named_modules():
        if id(m) in block_by_id:
            add_unit(name, list(m.parameters()), [m])
        elif id(m) not in block_ids:
            own = list(m.parameters(recurse=False))
            if own:
                add_unit(name or "root", own, [m])
    return units, owner
