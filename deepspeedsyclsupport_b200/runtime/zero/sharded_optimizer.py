"""ZeRO optimizer core shared by every stage (0-3), MiCS and hpZ-style sub-group sharding.

What the reference spreads over stage_1_and_2.py (2.5k lines), stage3.py (2.7k), parameter_offload.py and
partitioned_param_coordinator.py is here one class over the flat unit store (param_store.py):

  backward : gradients land in a contiguous per-unit buffer (autograd hook copy, or written there directly by the
             fused linear's wgrad GEMM -- `param.ds_get_main_grad()`); when a unit is complete its buffer is
             reduce-scattered on a side stream (stage >= 2) and accumulated into this rank's flat gradient shard.
             Reference: stage3.py:1110-1291 (__add_grad_to_ipg_bucket / __reduce_and_partition_ipg_grads),
             stage_1_and_2.py:860-1104 (hooks, average_tensor).
  step     : device-side overflow flag + global norm + clip coefficient (no host sync for bf16), then fused Adam over
             the flat fp32 shard writing the 16-bit shard in the same pass; stages 1/2 all-gather the units.
             Reference: stage3.py:1955-2017, stage_1_and_2.py:1771-1888.
  stage 3  : units are materialised around their module's forward/backward by hooks, with the fetch order recorded
             on the first iteration and replayed as prefetch on a side stream (coordinator.py:256-460).
  offload  : optimizer state (and fp32 master) in pinned host memory, AVX-512 Adam, `ratio` splits the flat shard
             between host and device steps (Offload++ / Twin-Flow, stage3.py:926-942).
"""
import math
import os
from collections import OrderedDict

import torch

from ... import comm as dist
from ...accelerator import get_accelerator
from ...ops import flat_optim
from ...utils.logging import log_dist, logger
from ..fp16.loss_scaler import CreateLossScaler
from .param_store import ParamUnit, ZeroParamStatus, build_units

_ADAM_TYPES = ("Adam", "AdamW", "FusedAdam", "DeepSpeedCPUAdam")
_LAST_OPT = None


def _in_backward():
    try:
        return torch._C._current_graph_task_id() != -1
    except Exception:
        return False


class _Region:
    """A contiguous slice [lo, hi) of the rank-flat shard whose optimizer state lives on one device."""

    def __init__(self, lo, hi, device, pin):
        self.lo, self.hi, self.device = lo, hi, device
        n = hi - lo
        kw = dict(dtype=torch.float32, device=device)
        self.master = torch.empty(n, **kw)
        self.state = {}
        self.pin = pin and device.type == "cpu" and torch.cuda.is_available()
        if self.pin:
            self.master = self.master.pin_memory()
        self.host_grad = None
        self.host_lp = None

    def zeros(self):
        t = torch.zeros(self.hi - self.lo, dtype=torch.float32, device=self.device)
        return t.pin_memory() if self.pin else t


class _NvmeRegion(_Region):
    """Optimizer state of [lo, hi) kept in files (ZeRO-Infinity): the fp32 master and the moments are cut into sub-groups of
    `sub_group` elements and streamed through two pinned buffers around each sub-group's update (reads of group g+1 overlap
    the update of g, writes of g overlap the update of g+1; runtime/swap_tensor/optimizer_swapper.py over csrc/host/aio.cpp).
    Everything else (fragment accessors, checkpoints, client optimizers) sees ordinary host tensors: touching `.master` /
    `.state` makes the region resident, the next step writes it back to the files and drops the host copy.
    (reference: stage3.py:561-595,926-942 + swap_tensor/partitioned_optimizer_swapper.py, pipelined_optimizer_swapper.py)"""
    is_nvme = True

    def __init__(self, lo, hi, folder, aio_config, sub_group, rank, pin):
        self.lo, self.hi, self.device = lo, hi, torch.device("cpu")
        self.pin = pin and torch.cuda.is_available()
        self.host_grad = self.host_lp = None
        n = hi - lo
        sub_group = max(1, min(int(sub_group), n))
        self.bounds = [(a, min(a + sub_group, n)) for a in range(0, n, sub_group)]
        self.folder, self.aio_config, self.rank = folder, dict(aio_config or {}), rank
        self._master = torch.empty(n, dtype=torch.float32)  # resident until the first step
        self._state = {}
        self.swapper = None

    # resident view (slow paths)
    def _materialize(self):
        if self._master is not None:
            return
        n = self.hi - self.lo
        names = self.swapper.state_names
        full = {k: torch.empty(n, dtype=torch.float32) for k in names}
        for g, (a, b) in enumerate(self.bounds):
            bufs = self.swapper.swap_in(g)
            for k in names:
                full[k][a:b].copy_(bufs[k])
        self._master = full.pop("master")
        self._state = full

    @property
    def master(self):
        self._materialize()
        return self._master

    @master.setter
    def master(self, v):
        self._master = v

    @property
    def state(self):
        self._materialize()
        return self._state

    @state.setter
    def state(self, v):
        self._materialize()
        self._state = v

    @property
    def resident(self):
        return self._master is not None

    def swap_out_all(self):
        """Resident tensors -> files (creating the swapper on first use); the host copy is dropped."""
        if self._master is None:
            return
        from ..swap_tensor.optimizer_swapper import OptimizerStateSwapper
        names = ("master", ) + tuple(sorted(self._state.keys()))
        if self.swapper is None or tuple(self.swapper.state_names) != names:
            self.swapper = OptimizerStateSwapper(self.folder, self.aio_config, [b - a for a, b in self.bounds], rank=self.rank,
                                                 state_names=names)
        for g, (a, b) in enumerate(self.bounds):
            tensors = {"master": self._master[a:b]}
            tensors.update({k: v[a:b] for k, v in self._state.items()})
            self.swapper.initialize(g, tensors)
        self._master, self._state = None, {}


class ZeroShardedOptimizer:

    def __init__(self, module, init_optimizer, *, stage, zero_config, dtype, dp_group=None, shard_group=None,
                 replica_group=None, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None,
                 clip_grad=0.0, gradient_accumulation_steps=1, grad_accum_dtype=None, communication_data_type=None,
                 mpu=None, timers=None, gradient_predivide_factor=1.0, prescale_gradients=False, grad_divisor=None,
                 broadcast_src=0, elastic_checkpoint=False, param_filter=None, aio_config=None):
        self.module = module
        self.aio_config = aio_config
        self.optimizer = init_optimizer
        self.stage = int(stage)
        self.zero_config = zero_config
        self.dtype = dtype
        self.mpu = mpu
        self.timers = timers
        self.clip_grad = float(clip_grad or 0.0)
        self.gas = gradient_accumulation_steps
        self.micro_step_id = 0
        self.overflow = False
        self.skipped_steps = 0
        self.custom_loss_scaler = False
        self.external_loss_scale = None
        self._global_grad_norm = None
        self.partition_gradients = self.stage >= 2
        self.elastic_checkpoint = elastic_checkpoint

        self.dp_group = dp_group
        if self.stage == 0:
            # optimizer states are replicated: shard world is 1 and gradients are all-reduced over the dp group
            self.shard_group, self.replica_group = None, dp_group
            self.shard_world, self.shard_rank = 1, 0
            self.replica_world = dist.get_world_size(dp_group) if dist.is_initialized() else 1
        else:
            self.shard_group = shard_group if shard_group is not None else dp_group
            self.replica_group = replica_group
            self.shard_world = dist.get_world_size(self.shard_group) if dist.is_initialized() else 1
            self.shard_rank = dist.get_rank(self.shard_group) if dist.is_initialized() else 0
            self.replica_world = dist.get_world_size(replica_group) if (replica_group is not None and dist.is_initialized()) else 1
        self.dp_world = dist.get_world_size(dp_group) if dist.is_initialized() else 1
        self.partition_count = self.shard_world
        # average over every rank that saw different data (sequence-parallel ranks share samples: engine passes divisor)
        self.grad_divisor = float(grad_divisor) if grad_divisor else float(self.shard_world * self.replica_world)
        self.predivide = float(gradient_predivide_factor) if prescale_gradients else 1.0

        acc = get_accelerator()
        self.device = torch.device(acc.current_device_name())
        self.cuda = self.device.type == "cuda"
        self.overlap_comm = bool(zero_config.overlap_comm) and self.cuda and self.shard_world * self.replica_world > 1
        self.comm_dtype = communication_data_type or dtype
        self.ag_stream = torch.cuda.Stream() if self.cuda else None
        self.rs_stream = torch.cuda.Stream() if self.cuda else None
        self.offload_stream = torch.cuda.Stream() if self.cuda else None
        self._d2h_done = []

        # ---- parameter groups -> units
        self.param_groups = self.optimizer.param_groups if self.optimizer is not None else []
        param_to_group = {}
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                param_to_group[id(p)] = gi
        for p in module.parameters():
            if id(p) not in param_to_group:
                param_to_group[id(p)] = -1  # frozen / not optimised
        self.units, self.owner = build_units(module, param_to_group, self.shard_world, self.shard_rank, dtype, self.device,
                                             param_filter=param_filter)
        self.persistence_threshold = int(zero_config.param_persistence_threshold) if self.stage == 3 else 0
        self._setup_param_attrs()
        self._allocate(broadcast_src)
        self._build_runs()

        if grad_accum_dtype is None:
            grad_accum_dtype = torch.float32 if (self.gas > 1 or dtype == torch.float32) else dtype
        self.grad_accum_dtype = grad_accum_dtype
        self._allocate_grads()

        self.loss_scaler = CreateLossScaler(dtype=dtype, static_loss_scale=static_loss_scale, dynamic_scaling=dynamic_loss_scale,
                                            dynamic_loss_args=dynamic_loss_args)
        self.dynamic_loss_scale = self.loss_scaler.dynamic
        self._inv_scale = torch.ones(1, dtype=torch.float32, device=self.device)
        self._skip_flag = torch.zeros(1, dtype=torch.int32, device=self.device)

        global _LAST_OPT
        _LAST_OPT = self  # debugging / test hook: the most recently built ZeRO optimizer of this process
        self._opt_kind = type(self.optimizer).__name__ if self.optimizer is not None else None
        self._fwd_keep = None
        self._local_grads_only = False
        self._step_count = 0
        self._register_grad_hooks()
        self._trace = {"fwd": [], "bwd": []}
        self._trace_frozen = False
        self._trace_pos = {"fwd": 0, "bwd": 0}
        if self._fetching:
            self._register_module_hooks()
        log_dist(f"ZeRO stage {self.stage}: {len(self.units)} units, shard world {self.shard_world}, replica world "
                 f"{self.replica_world}, {self.total_shard * self.shard_world / 1e6:.1f}M padded elements, "
                 f"shard {self.total_shard / 1e6:.1f}M, grad accum dtype {self.grad_accum_dtype}", ranks=[0])

    # ------------------------------------------------------------------ setup
    def _setup_param_attrs(self):
        pid = 0
        for u in self.units:
            for i, p in enumerate(u.params):
                if not getattr(p, "ds_init_sharded", False):
                    p.ds_shape = p.shape
                p.ds_numel = u.numels[i]
                p.ds_id = pid
                pid += 1
                p.ds_unit = u
                p.ds_index = i
                p.ds_status = ZeroParamStatus.AVAILABLE
                p.ds_persist = False
                p.ds_process_group = self.shard_group
                p.ds_zero_optimizer = self
                self._attach_param_methods(p)
            if self.stage == 3 and u.numel <= self.persistence_threshold:
                u.persistent = True
                for p in u.params:
                    p.ds_persist = True

    def _attach_param_methods(self, p):
        opt = self

        def all_gather(param_list=None, async_op=False, hierarchy=0):
            opt.fetch_units({id(q.ds_unit): q.ds_unit for q in (param_list or [p])}.values())

        def partition(param_list=None, hierarchy=0, has_been_updated=False):
            for u in {id(q.ds_unit): q.ds_unit for q in (param_list or [p])}.values():
                if has_been_updated:
                    opt._write_back_unit(u)
                opt.release_unit(u)

        def ds_summary(slf=p, use_debug_name=False):
            return {"id": slf.ds_id, "status": slf.ds_status.name, "numel": slf.numel(), "ds_numel": slf.ds_numel,
                    "shape": tuple(slf.shape), "ds_shape": tuple(slf.ds_shape), "requires_grad": slf.requires_grad,
                    "grad_shape": tuple(slf.grad.shape) if slf.grad is not None else None, "persist": slf.ds_persist,
                    "unit": slf.ds_unit.name, "ds_tensor.shape": tuple(slf.ds_tensor.shape)}

        p.all_gather = all_gather
        p.partition = partition
        p.ds_summary = ds_summary
        p.ds_get_main_grad = lambda q=p: opt.main_grad_view(q)
        p.ds_take_main_grad = lambda q=p: opt.main_grad_take(q)
        p.ds_main_grad_done = lambda q=p: opt._param_grad_ready(q)
        p.ds_grad_fresh = False

    def _unit_flat_from_params(self, u, broadcast_src):
        """Assemble the unit's full flat buffer from whatever form its parameters currently have."""
        flat = torch.zeros(u.padded, dtype=self.dtype, device=self.device)
        for i, p in enumerate(u.params):
            off, n = u.param_slice(i)
            if getattr(p, "ds_init_sharded", False):
                from .partition_parameters import gather_init_sharded
                flat[off:off + n].copy_(gather_init_sharded(p).view(-1))
            else:
                flat[off:off + n].copy_(p.data.reshape(-1).to(self.device, self.dtype))
        if dist.is_initialized() and broadcast_src is not None and self.dp_world > 1 and not all(
                getattr(p, "ds_init_sharded", False) for p in u.params):
            dist.broadcast(flat, src=dist.get_global_rank(self.dp_group, 0), group=self.dp_group)
        return flat

    def _allocate(self, broadcast_src):
        off = 0
        for u in self.units:
            u.shard_off = off
            off += u.shard_numel
        self.total_shard = off
        # ---- ZeRO-Infinity parameter tier: the 16-bit shards live in (pinned) host memory, or in a memory-mapped file for
        # device = nvme; a unit exists on the device only between its fetch and its release, also at shard world 1
        # (reference: offload_param, partition_parameters.py remote_device / partitioned_param_swapper.py)
        op = getattr(self.zero_config, "offload_param", None)
        op_dev = str(getattr(op.device, "value", op.device)) if op is not None else "none"
        self.offload_param = self.stage == 3 and op_dev in ("cpu", "nvme")
        self.p2p = None
        self._setup_p2p()
        if self.p2p is not None:
            self.lp_shard = self.p2p.view(self._p2p_lp_off, (self.total_shard, ), self.dtype)
            self.lp_shard.zero_()
        elif self.offload_param and op_dev == "nvme":
            assert op.nvme_path is not None, "offload_param.device = nvme needs offload_param.nvme_path"
            rank = dist.get_rank() if dist.is_initialized() else 0
            folder = os.path.join(str(op.nvme_path), "zero_stage_3", f"rank{rank}")
            os.makedirs(folder, exist_ok=True)
            self._lp_file = os.path.join(folder, "lp_params.swp")
            nbytes = max(self.total_shard, 1) * torch.empty(0, dtype=self.dtype).element_size()
            with open(self._lp_file, "wb") as f:
                f.truncate(nbytes)
            self.lp_shard = torch.from_file(self._lp_file, shared=True, size=max(self.total_shard, 1), dtype=self.dtype)[:self.total_shard]
        elif self.offload_param:
            self.lp_shard = torch.zeros(self.total_shard, dtype=self.dtype)
            if self.cuda and getattr(op, "pin_memory", False):
                self.lp_shard = self.lp_shard.pin_memory()
        else:
            self.lp_shard = torch.zeros(self.total_shard, dtype=self.dtype, device=self.device)
        self._fetching = self.stage == 3 and (self.shard_world > 1 or self.offload_param)
        keep_full = self.stage <= 2 or (self.shard_world == 1 and not self.offload_param)
        self.lp_full = None
        if keep_full and self.shard_world > 1:
            self.lp_full = torch.empty(sum(u.padded for u in self.units), dtype=self.dtype, device=self.device)
        foff = 0
        for u in self.units:
            u.shard = self.lp_shard[u.shard_off:u.shard_off + u.shard_numel]
            flat = self._unit_flat_from_params(u, broadcast_src)
            lo, hi = u.shard_range()
            u.shard.copy_(flat[lo:hi])
            if self.shard_world == 1 and keep_full:
                u.full = u.shard  # parameters alias the shard: gather/release are the identity
                u.bind_params()
            elif keep_full:
                u.full = self.lp_full[foff:foff + u.padded]
                u.full.copy_(flat)
                foff += u.padded
                u.bind_params()
            else:
                if u.persistent:
                    u.full = flat
                    u.bind_params()
                else:
                    u.full = None
                    u.unbind_params()
            for i, p in enumerate(u.params):
                p.ds_init_sharded = False
                a, n = u.param_slice(i)
                s_lo, s_hi = max(a, lo), min(a + n, hi)
                p.ds_tensor = u.shard[s_lo - lo:s_hi - lo] if s_lo < s_hi else u.shard[0:0]
            del flat
        self.keep_full = keep_full
        if self.p2p is not None:
            torch.cuda.synchronize()
            dist.barrier(group=self.shard_group)  # every rank's shard is in place before anyone pulls from it
        # ---- optimizer-state regions (host / device split = Offload++ ratio)
        oc = self.zero_config.offload_optimizer
        offload = oc is not None and str(getattr(oc.device, "value", oc.device)) != "none"
        self.offload_optimizer = offload
        self.regions = []
        if offload:
            ratio = float(oc.ratio)
            cut = int(self.total_shard * ratio) // 64 * 64 if ratio < 1.0 else self.total_shard
            if cut > 0:
                if str(getattr(oc.device, "value", oc.device)) == "nvme":
                    assert oc.nvme_path is not None, "offload_optimizer.device = nvme needs offload_optimizer.nvme_path"
                    rank = dist.get_rank() if dist.is_initialized() else 0
                    self.regions.append(_NvmeRegion(0, cut, os.path.join(str(oc.nvme_path), "zero_stage_3"), getattr(self, "aio_config", None),
                                                    self.zero_config.sub_group_size, rank, oc.pin_memory))
                else:
                    self.regions.append(_Region(0, cut, torch.device("cpu"), oc.pin_memory))
            if cut < self.total_shard:
                self.regions.append(_Region(cut, self.total_shard, self.device, False))
        else:
            self.regions.append(_Region(0, self.total_shard, self.device, False))
        for r in self.regions:
            if self.dtype == torch.float32 and r.device == self.lp_shard.device and not getattr(r, "is_nvme", False):
                r.master = self.lp_shard[r.lo:r.hi]  # fp32 training: the parameters ARE the master weights
            else:
                r.master.copy_(self.lp_shard[r.lo:r.hi].float())

    # ------------------------------------------------------------------ peer-memory collectives (own kernels, no NCCL)
    P2P_GRAD_SLOTS = 3

    def _setup_p2p(self):
        """Place this rank's parameter shard and a small ring of unit gradient buffers in symmetric (peer-mapped)
        memory so the ZeRO-3 collectives run as our own pull kernels over NVLink (csrc/comm/symm.cu):
        all-gather = every rank copies peers' shards; reduce-scatter = every rank sums its slice over peers' buffers,
        fused with 1/world scaling, dtype cast, accumulation into the gradient shard and the grad-norm partial."""
        want = (self.cuda and self.stage == 3 and self.shard_world > 1 and self.replica_world == 1 and not self.offload_param and
                bool(getattr(self.zero_config, "fused_collectives", True)) and self.dtype == torch.bfloat16 and
                os.environ.get("DSB200_DISABLE_P2P", "0") != "1" and self.shard_world <= 16)
        if not want:
            return
        try:
            from ...parallel.symm import SymmetricBuffer
            esz = 2
            max_unit = max(u.padded for u in self.units)
            self._p2p_lp_off = 0
            lp_bytes = (self.total_shard * esz + 255) // 256 * 256
            self._p2p_slot_bytes = (max_unit * esz + 255) // 256 * 256
            self._p2p_grad_off = lp_bytes
            total = lp_bytes + self.P2P_GRAD_SLOTS * self._p2p_slot_bytes
            # ZeRO++ regions (block-scaled fp8: 1 payload byte per element + one fp32 scale per 128 elements)
            self._qwz = bool(self.zero_config.zero_quantized_weights)
            self._qgz = bool(self.zero_config.zero_quantized_gradients)
            r256 = lambda n: (n + 255) // 256 * 256
            if self._qgz:  # per gradient slot: payload + scales of one whole unit
                self._qgz_q_off, self._qgz_q_bytes = total, r256(max_unit)
                total += self.P2P_GRAD_SLOTS * self._qgz_q_bytes
                self._qgz_s_off, self._qgz_s_bytes = total, r256(max_unit // 128 * 4)
                total += self.P2P_GRAD_SLOTS * self._qgz_s_bytes
            if self._qwz:  # quantised twin of this rank's parameter shard, refreshed after every step
                self._qwz_q_off = total
                total += r256(self.total_shard)
                self._qwz_s_off = total
                total += r256(self.total_shard // 128 * 4)
            self.p2p = SymmetricBuffer(self.shard_group, total)
            if self._qgz:
                self._qgz_q = [self.p2p.view(self._qgz_q_off + k * self._qgz_q_bytes, (max_unit, ), torch.uint8) for k in range(self.P2P_GRAD_SLOTS)]
                self._qgz_s = [self.p2p.view(self._qgz_s_off + k * self._qgz_s_bytes, (max_unit // 128, ), torch.float32) for k in range(self.P2P_GRAD_SLOTS)]
            if self._qwz:
                self._qwz_q = self.p2p.view(self._qwz_q_off, (self.total_shard, ), torch.uint8)
                self._qwz_s = self.p2p.view(self._qwz_s_off, (self.total_shard // 128, ), torch.float32)
            hpz = int(getattr(self.zero_config, "zero_hpz_partition_size", 1) or 1)
            if hpz > 1:
                # hpZ keeps a secondary shard inside the node so the backward all-gather stays off the inter-node links; this
                # shard group is one NVSwitch domain, where every peer is already "intra-node" at full bandwidth
                assert self.shard_world % hpz == 0 or hpz % self.shard_world == 0, \
                    f"zero_hpz_partition_size={hpz} must divide (or be a multiple of) the shard world {self.shard_world}"
                logger.info(f"ZeRO++ hpZ: partition size {hpz} inside one NVSwitch domain of {self.shard_world} GPUs -- the primary "
                            "shards already are the intra-node secondary shards; no second copy is kept")
            self._p2p_slot_views = [self.p2p.view(self._p2p_grad_off + k * self._p2p_slot_bytes, (max_unit, ), torch.bfloat16)
                                    for k in range(self.P2P_GRAD_SLOTS)]
            self._p2p_next_slot = 0
            self._p2p_reduce_count = 0
            self._p2p_barrier_events = {}
            self._p2p_sumsq = torch.zeros(1, dtype=torch.float32, device=self.device)
            self.nvlink_rx_bytes = 0  # running count of payload bytes this rank receives over NVLink (algorithmic, for reports)
            self._p2p_shards_synced = False
            self._p2p_ag_mode = os.environ.get("DSB200_AG_MODE", "ce")       # ce = copy engines, sm = pull kernel
            self._p2p_rs_mode = os.environ.get("DSB200_RS_MODE", "nvls" if self.p2p.mc_ptr else "pull")
            if self._p2p_rs_mode == "nvls" and not self.p2p.mc_ptr:
                self._p2p_rs_mode = "pull"
            self._p2p_comm_ctas = int(os.environ.get("DSB200_COMM_CTAS", "0"))
            if self._qgz:
                self._p2p_rs_mode = "fp8-pull"
            logger.info(f"ZeRO-3 peer-memory collectives: backend={self.p2p.backend} all-gather={self._p2p_ag_mode}"
                        f"{' (qwZ fp8 payload)' if self._qwz else ''} reduce-scatter={self._p2p_rs_mode}")
            self.comm_dtype = torch.bfloat16
        except Exception as e:  # pragma: no cover - peer access unavailable: keep the NCCL path
            logger.warning(f"peer-memory collectives unavailable ({e}); using NCCL")
            self.p2p = None

    def _p2p_take_slot(self, u):
        """Assign the next gradient slot of the ring to unit `u` (compute stream waits until every peer has finished
        reading the slot's previous contents: they have once this rank passed the barrier of reduce n-K+1)."""
        k = self._p2p_next_slot % self.P2P_GRAD_SLOTS
        n = self._p2p_next_slot
        self._p2p_next_slot += 1
        need = n - self.P2P_GRAD_SLOTS + 1
        if need >= 0 and need in self._p2p_barrier_events:
            t = self._probe_begin("grad_slot_wait")
            torch.cuda.current_stream().wait_event(self._p2p_barrier_events[need])
            self._probe_end(t)
        u.p2p_slot = k
        buf = self._p2p_slot_views[k][:u.padded]
        # the slot still holds another unit's gradients: parameters overwrite their slices on first touch (ds_grad_fresh),
        # only the gaps between them are cleared
        for off, n in self._unit_gaps(u):
            buf[off:off + n].zero_()
        self._mark_fresh([u])
        return buf

    def _p2p_reduce_unit(self, u):
        self._settle_fresh([u])
        stream = self.rs_stream
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            self.p2p.barrier()  # every rank has finished writing this unit's gradients (and reading older slots)
            ev = torch.cuda.Event()
            ev.record(stream)
            self._p2p_barrier_events[self._p2p_reduce_count] = ev
            self._p2p_barrier_events.pop(self._p2p_reduce_count - 2 * self.P2P_GRAD_SLOTS, None)
            self._p2p_reduce_count += 1
            dst = self.grad_shard[u.shard_off:u.shard_off + u.shard_numel]
            fuse_norm = self.gas == 1
            if self._rs_adam_enabled() and self._p2p_rs_adam_unit(u):
                u.full_grad = None
                u.reduced = True
                return
            if self._qgz and u.shard_numel % 128 == 0:
                self.nvlink_rx_bytes += (self.shard_world - 1) * (u.shard_numel + u.shard_numel // 128 * 4)
                return self._p2p_reduce_unit_qgz(u, dst, fuse_norm)
            # NVLS: the switch delivers the finished sum (1x the slice); pull: one read of the slice per peer
            self.nvlink_rx_bytes += u.shard_numel * 2 * (1 if self._p2p_rs_mode == "nvls" else self.shard_world - 1)
            rs = self.p2p.nvls_reduce_scatter if self._p2p_rs_mode == "nvls" else self.p2p.reduce_scatter
            rs(dst, self._p2p_grad_off + u.p2p_slot * self._p2p_slot_bytes, torch.bfloat16, u.shard_numel,
               scale=1.0 / self.grad_divisor, accumulate=self.gas > 1, sumsq=self._p2p_sumsq if fuse_norm else None,
               max_ctas=self._p2p_comm_ctas)
        u.full_grad = None
        u.reduced = True
        self._offload_grad_async(u, stream)

    def _p2p_reduce_unit_qgz(self, u, dst, fuse_norm):
        """qgZ on one NVSwitch hop: quantise this rank's whole unit gradient to block-scaled fp8 inside its symmetric slot,
        barrier, then every rank pulls ITS slice of payload + scales from all peers and dequant-reduces (fp32 sum) straight
        into the gradient shard, with scale / sum-of-squares fused.  (Called on the reduce stream; the bf16 -> fp8 barrier
        replaces the plain one.)"""
        from ...ops import quantizer as Q
        k = u.p2p_slot
        Q.quantize_fp8_block(self._p2p_slot_views[k][:u.padded], out=(self._qgz_q[k][:u.padded], self._qgz_s[k][:u.padded // 128]))
        self.p2p.barrier()  # every rank's payload is written
        self.p2p.fp8_reduce_scatter(dst, self._qgz_q_off + k * self._qgz_q_bytes, self._qgz_s_off + k * self._qgz_s_bytes,
                                    u.shard_numel, scale=1.0 / self.grad_divisor, accumulate=self.gas > 1,
                                    sumsq=self._p2p_sumsq if fuse_norm else None, max_ctas=self._p2p_comm_ctas)
        u.full_grad = None
        u.reduced = True
        self._zeropp_serialise(self.rs_stream)

    # ---- reduce-scatter fused with the partitioned AdamW update (csrc/comm/symm.cu rs_adam_kernel)
    def _rs_adam_enabled(self):
        """The update may run inside backward, unit by unit, only when nothing global has to be known first: no gradient
        clipping (the clip factor needs the norm of ALL gradients), no accumulation, no loss scaling / overflow skip (bf16),
        a fused Adam flavour, optimizer state resident on the device.  With clipping the step stays two-phase: the fused
        reduce-scatter produces shard + sum of squares during backward, then one Adam pass reads the clip factor on the device."""
        ok = getattr(self, "_rs_adam_ok", None)
        if ok is None:
            ok = (self.p2p is not None and self.gas == 1 and self.clip_grad == 0.0 and self.dtype == torch.bfloat16 and
                  self._opt_kind in _ADAM_TYPES and not self.custom_loss_scaler and len(self.regions) == 1 and
                  self.regions[0].device.type == "cuda" and not getattr(self.regions[0], "is_nvme", False) and
                  self.replica_world == 1 and os.environ.get("DSB200_FUSED_RS_ADAM", "1") != "0")
            self._rs_adam_ok = ok
            if ok:
                logger.info("ZeRO-3: gradient reduce-scatter fused with the partitioned AdamW update (no clipping, gas == 1)")
        return ok

    def _unit_runs(self, u):
        """[(group, lo, hi)] in rank-flat shard coordinates for one unit (trainable ranges only)."""
        runs = getattr(u, "_flat_runs", None)
        if runs is None:
            runs = u._flat_runs = [(g, u.shard_off + a, u.shard_off + b) for g, a, b in u.shard_group_runs() if g >= 0 and b > a]
        return runs

    def _p2p_rs_adam_unit(self, u):
        runs = self._unit_runs(u)
        if any((lo - u.shard_off) % 8 or (hi - lo) % 8 for _, lo, hi in runs):
            return False
        region = self.regions[0]
        self._ensure_state(region, self._opt_kind)
        kind = self._opt_kind
        adamw = False if kind == "Adam" else (True if kind == "AdamW" else bool(getattr(self.optimizer, "adam_w_mode", True)))
        g_out = self.grad_shard if self.grad_shard.dtype == torch.bfloat16 else None
        src_off = self._p2p_grad_off + u.p2p_slot * self._p2p_slot_bytes
        for gi, lo, hi in runs:
            hp = self.param_groups[gi]
            b1, b2 = hp.get("betas", (0.9, 0.999))
            self.p2p.rs_adam_range(src_off, u.shard_numel, lo - u.shard_off, region.master[lo:hi], region.state["exp_avg"][lo:hi],
                                   region.state["exp_avg_sq"][lo:hi], self.lp_shard[lo:hi], scale=1.0 / self.grad_divisor,
                                   lr=hp["lr"], beta1=b1, beta2=b2, eps=hp.get("eps", 1e-8), weight_decay=hp.get("weight_decay", 0.0),
                                   adamw_mode=adamw, bias_correction=hp.get("bias_correction", True), step=self._step_count + 1,
                                   g_out=g_out[lo:hi] if g_out is not None else None, sumsq=self._p2p_sumsq,
                                   nvls=self._p2p_rs_mode == "nvls", max_ctas=self._p2p_comm_ctas)
        u.adam_done = True
        return True

    def _build_runs(self):
        """(group_id, lo, hi) runs over the rank-flat shard space, merged across unit boundaries."""
        runs = []
        for u in self.units:
            for g, a, b in u.shard_group_runs():
                lo, hi = u.shard_off + a, u.shard_off + b
                if runs and runs[-1][0] == g and runs[-1][2] == lo:
                    runs[-1][2] = hi
                else:
                    runs.append([g, lo, hi])
        self.runs = [tuple(r) for r in runs]

    def _allocate_grads(self):
        self.grad_shard = torch.zeros(self.total_shard, dtype=self.grad_accum_dtype, device=self.device)
        self.grad_full_persistent = None
        self._grad_is_shard = self.shard_world == 1
        if self.shard_world == 1:
            # full gradient == shard: hooks / fused wgrad write straight into the accumulation buffer
            for u in self.units:
                u.full_grad = self.grad_shard[u.shard_off:u.shard_off + u.shard_numel]
        elif self.stage <= 1:
            self.grad_full_persistent = torch.zeros(sum(u.padded for u in self.units), dtype=self.grad_accum_dtype,
                                                    device=self.device)
            off = 0
            for u in self.units:
                u.full_grad = self.grad_full_persistent[off:off + u.padded]
                off += u.padded
        self._grads_dirty = False

    # ------------------------------------------------------------------ gradient path
    def _register_grad_hooks(self):
        self._hook_handles = []
        for u in self.units:
            for p in u.trainable:
                self._hook_handles.append(p.register_post_accumulate_grad_hook(self._make_grad_hook(p)))

    def _make_grad_hook(self, p):

        def hook(param):
            if param.grad is None:
                return
            view = self.main_grad_view(param)
            g = param.grad
            if view.dtype == g.dtype and g.shape == view.shape:
                view.add_(g)
            else:
                view.add_(g.to(view.dtype).view_as(view))
            param.grad = None
            self._param_grad_ready(param)

        return hook

    def _grad_slice(self, p):
        u = p.ds_unit
        if u.full_grad is None:
            if self.p2p is not None:
                u.full_grad = self._p2p_take_slot(u)
            else:
                u.full_grad = torch.zeros(u.padded, dtype=self.comm_dtype, device=self.device)
        off, n = u.param_slice(p.ds_index)
        self._grads_dirty = True
        return u.full_grad[off:off + n].view(p.ds_shape)

    def main_grad_view(self, p):
        """Slice of the unit's contiguous gradient buffer for `p` (allocated on first use in a backward pass), to be
        ACCUMULATED into.  Buffers are not memset between steps: a parameter whose slice still holds stale data is
        flagged `ds_grad_fresh` and zeroed here, on first touch."""
        view = self._grad_slice(p)
        if p.ds_grad_fresh:
            view.zero_()
            p.ds_grad_fresh = False
        return view

    def main_grad_take(self, p):
        """-> (slice, fresh).  fresh=True: the caller must OVERWRITE the slice (first gradient of this parameter since the
        buffer was recycled) -- the fused wgrad GEMM then stores instead of reduce-adding, and nobody zero-fills."""
        view = self._grad_slice(p)
        fresh = p.ds_grad_fresh
        p.ds_grad_fresh = False
        return view, fresh

    def _mark_fresh(self, units):
        for u in units:
            for q in u.trainable:
                q.ds_grad_fresh = True

    def _settle_fresh(self, units):
        """Parameters that received no gradient keep stale data in their slice: zero them before the buffer is consumed."""
        for u in units:
            for q in u.trainable:
                if q.ds_grad_fresh:
                    if u.full_grad is not None:
                        off, n = u.param_slice(q.ds_index)
                        u.full_grad[off:off + n].zero_()
                    q.ds_grad_fresh = False

    def _unit_gaps(self, u):
        """[(off, len)] of the unit-flat space not covered by a trainable parameter (alignment padding, frozen weights)."""
        gaps = getattr(u, "_grad_gaps", None)
        if gaps is None:
            spans = sorted(u.param_slice(q.ds_index) for q in u.trainable)
            gaps, cur = [], 0
            for off, n in spans:
                if off > cur:
                    gaps.append((cur, off - cur))
                cur = max(cur, off + n)
            if cur < u.padded:
                gaps.append((cur, u.padded - cur))
            u._grad_gaps = gaps
        return gaps

    def _param_grad_ready(self, p):
        u = p.ds_unit
        u.ready += 1
        if u.ready >= len(u.trainable):
            self._unit_grads_done(u)

    def _unit_grads_done(self, u):
        u.ready = 0
        if self.shard_world == 1 and self.replica_world == 1:
            u.reduced = True
            if self.cuda and self._grad_is_shard:
                self._offload_grad_async(u, torch.cuda.current_stream())
        elif self.stage >= 2 and self.shard_world > 1:
            self._reduce_unit(u)
        # stage 0/1: reduction happens once at the accumulation boundary (reduce_gradients)
        if self._fetching and not u.persistent and _in_backward():
            self.release_unit(u)

    def _reduce_unit(self, u):
        """reduce-scatter (+ replica all-reduce) of one unit's gradient buffer into the flat gradient shard."""
        if u.full_grad is None:
            return
        if self._local_grads_only:
            # communication-compressed optimizers (1-bit Adam & co.) consume the LOCAL gradient after their warm-up: no
            # collective, no averaging (stage 0 only, where the "shard" is the whole unit)
            assert self.shard_world == 1, "compressed-communication optimizers require ZeRO stage 0"
            dst = self.grad_shard[u.shard_off:u.shard_off + u.shard_numel]
            if u.full_grad.data_ptr() != dst.data_ptr():
                dst.add_(u.full_grad.to(dst.dtype) if u.full_grad.dtype != dst.dtype else u.full_grad)
            u.reduced = True
            return
        if self.p2p is not None:
            return self._p2p_reduce_unit(u)
        buf = u.full_grad
        stream = self.rs_stream if self.overlap_comm else None
        if stream is not None:
            stream.wait_stream(torch.cuda.current_stream())
        ctx = torch.cuda.stream(stream) if stream is not None else _nullctx()
        with ctx:
            if self.predivide != 1.0:
                buf.div_(self.predivide)
            dst = self.grad_shard[u.shard_off:u.shard_off + u.shard_numel]
            if self.shard_world > 1:
                tmp = torch.empty(u.shard_numel, dtype=buf.dtype, device=self.device)
                dist.reduce_scatter_tensor(tmp, buf, op=dist.ReduceOp.SUM, group=self.shard_group)
            else:
                tmp = buf
            if self.replica_world > 1:
                dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=self.replica_group)
            scale = self.predivide / self.grad_divisor
            if tmp.data_ptr() != dst.data_ptr():
                dst.add_(tmp.to(dst.dtype) if tmp.dtype != dst.dtype else tmp, alpha=scale)
            elif scale != 1.0:
                dst.mul_(scale)
            if stream is not None:
                buf.record_stream(stream)
                tmp.record_stream(stream)
        if self.grad_full_persistent is None and self.shard_world > 1:
            u.full_grad = None
        u.reduced = True
        if self.cuda:
            self._offload_grad_async(u, stream if stream is not None else torch.cuda.current_stream())

    def reduce_gradients(self, pipeline_parallel=False):
        """Called by the engine after backward at an accumulation boundary (stage 0/1) or every micro-step (stage >= 2)
        to flush units whose gradients did not all arrive (unused parameters)."""
        boundary_only = self.stage <= 1
        for u in self.units:
            if self.shard_world == 1 and self.replica_world == 1:
                continue
            if boundary_only:
                self._reduce_unit(u)
                if self.grad_full_persistent is not None:
                    pass
            elif not u.reduced and (u.full_grad is not None and self.grad_full_persistent is None):
                self._reduce_unit(u)
            elif self.shard_world == 1 and self.replica_world > 1 and not u.reduced:
                self._reduce_unit(u)
        if self.overlap_comm:
            t = self._probe_begin("reduce_drain")
            torch.cuda.current_stream().wait_stream(self.rs_stream)
            self._probe_end(t)
        if boundary_only and self.grad_full_persistent is not None:
            self.grad_full_persistent.zero_()

    def overlapping_partition_gradients_reduce_epilogue(self):
        self.reduce_gradients()

    def end_micro_step(self, boundary, local_only=False):
        """Engine hook after every backward().  local_only: skip the data-parallel reduction (the basic optimizer
        communicates by itself, e.g. 1-bit Adam after its warm-up)."""
        self._local_grads_only = bool(local_only)
        if self._grad_is_shard:
            self._settle_fresh(self.units)
        if self.stage >= 2 or boundary:
            if self.stage <= 1 and not boundary:
                pass
            else:
                self.reduce_gradients()
        for u in self.units:
            u.reduced = False
            u.ready = 0
        if self._fetching:
            for u in self.units:
                if not u.persistent and u.status == ZeroParamStatus.AVAILABLE:
                    self.release_unit(u)
            self._end_of_pass()
        self.micro_step_id += 1

    # ------------------------------------------------------------------ stage-3 fetch / release
    def fetch_units(self, units, prefetch_phase=None):
        for u in units:
            self._fetch(u)
        if prefetch_phase is not None and self._trace_frozen:
            self._prefetch_next(prefetch_phase)

    def _launch_gather(self, u):
        if u.status != ZeroParamStatus.NOT_AVAILABLE:
            return
        stream = self.ag_stream if self.cuda else None
        ctx = torch.cuda.stream(stream) if stream is not None else _nullctx()
        main = torch.cuda.current_stream() if self.cuda else None
        with ctx:
            u.full = torch.empty(u.padded, dtype=self.dtype, device=self.device)
            if self.p2p is not None:
                self._p2p_sync_shards(stream, main)
                self._p2p_gather_unit(u)
            elif self.zero_config.zero_quantized_weights and self.cuda:
                from .zeropp import quantized_all_gather
                quantized_all_gather(u.full, u.shard.to(self.device, non_blocking=True) if self.offload_param else u.shard, self.shard_group)
            else:
                self._gather_unit_into(u.full, u)
            if stream is not None:
                u.gather_event = torch.cuda.Event()
                u.gather_event.record(stream)
        u.status = ZeroParamStatus.INFLIGHT

    P2P_CE_MIN_BYTES = 1 << 20

    def _p2p_gather_unit(self, u):
        """Pull every rank's shard of `u` into u.full.  Large shards go over the copy engines (NVLink DMA, no SMs: the
        persistent GEMMs keep all 148 SMs); small ones (norm weights) use one pull kernel, cheaper than W tiny DMAs."""
        nbytes = u.shard_numel * 2
        off = self._p2p_lp_off + u.shard_off * 2
        if self._qwz and nbytes >= self.P2P_CE_MIN_BYTES:
            self.nvlink_rx_bytes += (self.shard_world - 1) * (u.shard_numel + u.shard_numel // 128 * 4)
            return self._p2p_gather_unit_qwz(u)
        self.nvlink_rx_bytes += (self.shard_world - 1) * nbytes
        if nbytes >= self.P2P_CE_MIN_BYTES and self._p2p_ag_mode == "ce":
            self.p2p.all_gather_ce(u.full, off, nbytes)
        else:
            self.p2p.all_gather(u.full, off, nbytes, max_ctas=self._p2p_comm_ctas)

    def _p2p_gather_unit_qwz(self, u):
        """qwZ: peers' shards arrive as block-scaled fp8 (half the NVLink bytes of bf16) on the copy engines into a staging
        buffer; one local kernel dequantises into u.full.  This rank's own slice is copied in bf16 (exact)."""
        from ...ops import quantizer as Q
        n, W, r = u.shard_numel, self.shard_world, self.shard_rank
        stage_q = torch.empty(W * n, dtype=torch.uint8, device=self.device)
        stage_s = torch.empty(W * n // 128, dtype=torch.float32, device=self.device)
        self.p2p.all_gather_ce(stage_q, self._qwz_q_off + u.shard_off, n)
        self.p2p.all_gather_ce(stage_s, self._qwz_s_off + u.shard_off // 128 * 4, n // 128 * 4)
        Q.dequantize_fp8_block(stage_q, stage_s, out=u.full)
        u.full[r * n:(r + 1) * n].copy_(u.shard)
        stage_q.record_stream(torch.cuda.current_stream())
        stage_s.record_stream(torch.cuda.current_stream())
        self._zeropp_serialise(torch.cuda.current_stream())

    def _zeropp_serialise(self, stream):
        """UNVALIDATED mitigation for more than 4 ranks.  The ZeRO++ peer path runs at 2 and 4 GPUs; at 8 GPUs the asynchronous
        run dead-locks in a device barrier while the same run with CUDA_LAUNCH_BLOCKING=1 completes (profiles/zeropp_8gpu/), i.e.
        the hazard needs the host to run ahead of the device.  Until it is found, the host waits for every quantised gather /
        reduce chain on such groups, which bounds the run-ahead to one unit (at some cost in overlap).  DSB200_ZEROPP_UNIT_SYNC=0
        restores the fully asynchronous behaviour."""
        if self.shard_world > 4 and os.environ.get("DSB200_ZEROPP_UNIT_SYNC", "1") != "0":
            stream.synchronize()

    def _p2p_refresh_qwz(self):
        """Re-quantise this rank's parameter shard into its symmetric fp8 twin (after every update of the shard)."""
        from ...ops import quantizer as Q
        Q.quantize_fp8_block(self.lp_shard, out=(self._qwz_q, self._qwz_s))

    def _p2p_shards_written(self):
        """Call after writing this rank's 16-bit shard outside step() (write-back of gathered parameters, checkpoint load,
        safe_set_*): peers pull shards with one-sided loads, so the next gather must first pass a cross-rank barrier that is
        stream-ordered after the write.  The write sites are collective by contract, so every rank re-arms together."""
        self._p2p_shards_synced = False

    def _p2p_sync_shards(self, stream, main):
        """Device barrier (once per pass, and after any out-of-step shard write) before the first one-sided pull: a faster
        rank must not read a shard its owner is still writing.  Runs on the gather stream, ordered after everything the
        compute stream `main` has enqueued (the shard writes)."""
        if self._p2p_shards_synced:
            return
        if stream is not None and main is not None:
            stream.wait_stream(main)
        if self._qwz:
            self._p2p_refresh_qwz()
        self.p2p.barrier()
        self._p2p_shards_synced = True

    def _gather_unit_into(self, full, u):
        """full[padded] <- every rank's shard of the unit (host-resident shards are staged to the device first)."""
        shard = u.shard
        if self.shard_world == 1:
            full[:u.shard_numel].copy_(shard, non_blocking=True)
            return
        if shard.device != full.device:
            shard = shard.to(full.device, non_blocking=True)
        dist.all_gather_into_tensor(full, shard, group=self.shard_group)

    def _fetch(self, u):
        if u.status == ZeroParamStatus.AVAILABLE:
            return
        if u.status == ZeroParamStatus.NOT_AVAILABLE:
            self._launch_gather(u)
        if u.gather_event is not None:
            t = self._probe_begin("gather_wait")
            torch.cuda.current_stream().wait_event(u.gather_event)
            self._probe_end(t)
            u.full.record_stream(torch.cuda.current_stream())
            u.gather_event = None
        u.bind_params()

    # ---- exposed-communication probe: CUDA-event pairs around every point where the compute stream waits for, or itself
    # runs, communication.  Off by default (no events recorded); bench.py turns it on for a separate measurement phase.
    _probe = None

    def comm_probe_start(self):
        self._probe = []

    def comm_probe_stop(self, detail=False):
        """-> total milliseconds the compute stream spent stalled on / executing communication since comm_probe_start()
        (detail=True: also {site: ms} for gather_wait / grad_slot_wait / reduce_drain / grad_norm / post_step)."""
        pairs, self._probe = self._probe or [], None
        if self.cuda:
            torch.cuda.synchronize()
        by_site = {}
        self.comm_probe_trace = []  # [(site, ms)] in program order, for timelines
        for site, a, b in pairs:
            ms = a.elapsed_time(b)
            by_site[site] = by_site.get(site, 0.0) + ms
            self.comm_probe_trace.append((site, round(ms, 3)))
        total = float(sum(by_site.values()))
        return (total, by_site) if detail else total

    def _probe_begin(self, site="other"):
        if self._probe is None or not self.cuda:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return site, e

    def _probe_end(self, t):
        if t is None:
            return
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        self._probe.append((t[0], t[1], e1))

    def release_unit(self, u):
        if self.keep_full or u.persistent or u.status == ZeroParamStatus.NOT_AVAILABLE:
            return
        if u.status == ZeroParamStatus.INFLIGHT:
            self._fetch(u)
        u.unbind_params()
        u.full = None

    def _write_back_unit(self, u):
        """Copy this rank's slice of a (modified) full buffer back into the shard + fp32 master."""
        lo, hi = u.shard_range()
        if u.full is not None and u.full.data_ptr() != u.shard.data_ptr():
            u.shard.copy_(u.full[lo:hi])
            if self.p2p is not None:
                self._p2p_shards_written()
        self._refresh_master(u.shard_off, u.shard_off + u.shard_numel)

    def _refresh_master(self, lo, hi):
        for r in self.regions:
            a, b = max(lo, r.lo), min(hi, r.hi)
            if a < b:
                r.master[a - r.lo:b - r.lo].copy_(self.lp_shard[a:b].float())

    def _record(self, phase, u):
        if not self._trace_frozen:
            self._trace[phase].append(u.uid)

    def _prefetch_next(self, phase):
        order = self._trace[phase]
        pos = self._trace_pos[phase]
        budget = int(self.zero_config.prefetch_bucket_size)
        live = sum(x.padded for x in self.units if x.status != ZeroParamStatus.NOT_AVAILABLE and not x.persistent)
        k = pos
        while k < len(order) and budget > 0:
            nu = self.units[order[k]]
            if nu.status == ZeroParamStatus.NOT_AVAILABLE and not nu.persistent:
                if live + nu.padded > int(self.zero_config.max_live_parameters) and live > 0:
                    break
                self._launch_gather(nu)
                budget -= nu.padded
                live += nu.padded
            k += 1

    def _keep_after_forward(self, u):
        """ZeRO-3 releases a unit after its forward and gathers it again for backward.  Units whose gathered size, summed
        from their position to the END of the forward order (= the units backward needs first), fits in
        `stage3_max_live_parameters` stay resident instead: on a 180 GB GPU a 7B model keeps everything and the backward
        all-gathers disappear.  (Reference semantics: max_live_parameters / max_reuse_distance in
        partitioned_param_coordinator.py:262-270,428-433.)"""
        if not self._trace_frozen:
            return False
        if self._fwd_keep is None:
            budget = int(self.zero_config.max_live_parameters)
            keep, seen, total = {}, set(), 0
            for uid in reversed(self._trace["fwd"]):
                if uid in seen:
                    continue
                seen.add(uid)
                nu = self.units[uid]
                if not nu.persistent:
                    total += nu.padded
                keep[uid] = total <= budget
            self._fwd_keep = keep
        return self._fwd_keep.get(u.uid, False)

    def _end_of_pass(self):
        if self.p2p is not None:
            self._p2p_shards_synced = False  # the next pass (or step()'s own barrier) re-synchronises before any pull
        if not self._trace_frozen and self._trace["fwd"]:
            self._trace_frozen = True
        self._trace_pos = {"fwd": 0, "bwd": 0}

    def _register_module_hooks(self):
        mod_units = OrderedDict()
        for u in self.units:
            for m in u.modules:
                mod_units.setdefault(id(m), (m, []))[1].append(u)
        # modules that *use* a tied parameter owned by another unit must fetch that unit as well.  Only the parameters a hooked
        # module holds itself (directly, or through children without a hook of their own) count: descending into hooked
        # children would make the root module fetch -- and wait for -- every unit of the model before its first kernel
        def held(m, top=True):
            if not top and id(m) in mod_units:
                return
            yield from m.parameters(recurse=False)
            for child in m.children():
                yield from held(child, False)

        for m in self.module.modules():
            if id(m) in mod_units:
                for p in held(m):
                    u = self.owner.get(id(p))
                    if u is not None and u not in mod_units[id(m)][1]:
                        mod_units[id(m)][1].append(u)
        self._module_hooks = []
        for _, (m, units) in mod_units.items():
            self._hook_module(m, units)

    def _advance(self, phase, units):
        order = self._trace[phase]
        pos = self._trace_pos[phase]
        if self._trace_frozen:
            uids = [u.uid for u in units]
            while pos < len(order) and order[pos] not in uids:
                pos += 1
            self._trace_pos[phase] = min(pos + len(uids), len(order))

    def _hook_module(self, m, units):
        opt = self

        def pre_fwd(mod, args):
            phase = "bwd" if _in_backward() else "fwd"
            for u in units:
                opt._record(phase, u)
            opt._advance(phase, units)
            opt.fetch_units(units, prefetch_phase=phase)

        def post_fwd(mod, args, out):
            if _in_backward():
                return  # recompute inside backward: parameters are still needed for dgrad/wgrad
            keep_ok = mod.training and torch.is_grad_enabled()
            for u in units:
                if keep_ok and opt._keep_after_forward(u):
                    continue  # stays gathered until its backward: no second all-gather (stage3_max_live_parameters)
                opt.release_unit(u)

        def pre_bwd(mod, grad_out):
            for u in units:
                opt._record("bwd", u)
            opt._advance("bwd", units)
            opt.fetch_units(units, prefetch_phase="bwd")

        def post_bwd(mod, gin, gout):
            for u in units:
                if not u.trainable:
                    opt.release_unit(u)

        self._module_hooks.append(m.register_forward_pre_hook(pre_fwd))
        self._module_hooks.append(m.register_forward_hook(post_fwd))
        self._module_hooks.append(m.register_full_backward_pre_hook(pre_bwd))
        self._module_hooks.append(m.register_full_backward_hook(post_bwd))

    # ------------------------------------------------------------------ loss scaling / backward
    @property
    def loss_scale(self):
        if self.custom_loss_scaler:
            return self.external_loss_scale
        return self.loss_scaler.cur_scale

    cur_scale = loss_scale

    def backward(self, loss, retain_graph=False):
        if self.custom_loss_scaler:
            (self.external_loss_scale * loss).backward(retain_graph=retain_graph)
        else:
            self.loss_scaler.backward(loss.float(), retain_graph=retain_graph)

    def zero_grad(self, set_to_none=True):
        self._d2h_done = []  # early gradient copies to the host belong to the step that just ended (or was skipped)
        for u in self.units:
            for p in u.params:
                p.grad = None
        if self._grad_is_shard:
            # gradients are written straight into the accumulation buffer: mark instead of a 2-bytes-per-parameter memset;
            # the first wgrad of every parameter stores, untouched parameters are zeroed in end_micro_step
            self._mark_fresh(self.units)
        else:
            self.grad_shard.zero_()
            if self.grad_full_persistent is not None:
                self.grad_full_persistent.zero_()
        self._grads_dirty = False

    # ------------------------------------------------------------------ step
    def _model_parallel_all_reduce(self, tensor, op):
        if self.mpu is not None and hasattr(self.mpu, "get_model_parallel_group"):
            g = self.mpu.get_model_parallel_group()
            if g is not None and dist.get_world_size(g) > 1:
                dist.all_reduce(tensor, op=op, group=g)

    def _check_overflow(self):
        if self.dtype != torch.float16:
            return False
        if self.grad_shard.is_cuda:
            from ...ops.fused_ops import sumsq
            probe = sumsq(self.grad_shard)  # one fused pass; an inf / nan anywhere makes the sum of squares non-finite
        else:
            probe = self.grad_shard.float().sum().reshape(1)
        bad = (~torch.isfinite(probe)).to(torch.float32).reshape(1)
        if dist.is_initialized():
            if self.shard_world > 1:
                dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=self.shard_group)
            self._model_parallel_all_reduce(bad, dist.ReduceOp.MAX)
        return bool(bad.item() > 0)

    def _compute_global_norm(self):
        """Device-side ||g||_2 of the (still loss-scaled) gradient over all shards."""
        g = self.grad_shard
        if self.p2p is not None and self.gas == 1:
            # sum(g^2) was accumulated by the fused reduce-scatter kernels while the gradients were being reduced
            sq = self._p2p_sumsq.clone()
            self._p2p_sumsq.zero_()
            dist.all_reduce(sq, op=dist.ReduceOp.SUM, group=self.shard_group)
            self._model_parallel_all_reduce(sq, dist.ReduceOp.SUM)
            return sq.sqrt()
        if g.is_cuda:
            from ...ops.fused_ops import sumsq
            sq = sumsq(g)  # one fused pass over the flat gradient shard
        else:
            sq = (g.float() if g.dtype != torch.float32 else g).square().sum().reshape(1)
        if dist.is_initialized():
            if self.shard_world > 1:
                dist.all_reduce(sq, op=dist.ReduceOp.SUM, group=self.shard_group)
            self._model_parallel_all_reduce(sq, dist.ReduceOp.SUM)
        return sq.sqrt()

    def step(self, closure=None):
        if self.begin_step():
            return
        t = self._probe_begin("grad_norm") if self.shard_world * self.replica_world > 1 else None
        norm = self._compute_global_norm()
        self._probe_end(t)
        self.finish_step(norm)

    def begin_step(self):
        """Overflow handling + loss-scale update.  Returns True when the step must be skipped."""
        self.micro_step_id = 0
        if self.overlap_comm or self.p2p is not None:
            t = self._probe_begin("reduce_drain")
            torch.cuda.current_stream().wait_stream(self.rs_stream)
            self._probe_end(t)
        self.overflow = self._check_overflow()
        self._prev_scale = self.loss_scale
        if not self.custom_loss_scaler:
            self.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            self.skipped_steps += 1
            self.zero_grad()
            return True
        return False

    def finish_step(self, norm):
        """`norm`: device tensor with the (loss-scaled) global gradient norm over everything that is clipped together."""
        loss_scale = self._prev_scale if not self.custom_loss_scaler else self.external_loss_scale
        self._global_grad_norm = norm / loss_scale
        inv = torch.full_like(norm, 1.0 / loss_scale)
        if self.clip_grad > 0.0:
            clip = (self._global_grad_norm / self.clip_grad).clamp(min=1.0)
            inv = inv / clip
        self._inv_scale.copy_(inv)
        self._step_count += 1
        self._optimizer_step()
        t = self._probe_begin("post_step") if self.shard_world * self.replica_world > 1 else None
        self._post_step()
        self._probe_end(t)
        self.zero_grad()

    def _hyper(self, gi):
        g = self.param_groups[gi]
        return g

    def _ensure_state(self, region, kind):
        if region.state:
            return
        if kind in _ADAM_TYPES or kind in ("FusedLamb", "OnebitAdam", "ZeroOneAdam", "OnebitLamb"):
            region.state = {"exp_avg": region.zeros(), "exp_avg_sq": region.zeros()}
        elif kind in ("FusedLion", "DeepSpeedCPULion", "Lion"):
            region.state = {"exp_avg": region.zeros()}
        elif kind in ("Adagrad", "DeepSpeedCPUAdagrad"):
            region.state = {"exp_avg_sq": region.zeros()}

    def _is_fused_kind(self, kind):
        return kind in _ADAM_TYPES or kind in ("FusedLion", "DeepSpeedCPULion", "Lion", "Adagrad", "DeepSpeedCPUAdagrad")

    def _rebind_basic_optimizer(self):
        """Client / non-fused optimizers (SGD, LAMB, 1-bit Adam, anything torch.optim): point the user's optimizer object
        at fp32 views of the flat master regions -- one big tensor per (region, param group) -- and keep calling ITS step().
        State, hyper-parameters, LR schedulers and custom communication (compressed all-reduce) all stay with that
        object, which is what the reference's FP16/BF16/ZeRO wrappers do with their flattened fp32 groups."""
        self._generic_bind = []
        per_group = [[] for _ in self.param_groups]
        for region in self.regions:
            for gi, lo, hi in self.runs:
                a, b = max(lo, region.lo), min(hi, region.hi)
                if a >= b or gi < 0:
                    continue
                s = slice(a - region.lo, b - region.lo)
                w = torch.nn.Parameter(region.master[s], requires_grad=True)
                per_group[gi].append(w)
                self._generic_bind.append((w, region, s))
        self.optimizer.state.clear()
        for g, ws in zip(self.optimizer.param_groups, per_group):
            g["params"] = ws

    def _generic_step(self):
        if not hasattr(self, "_generic_bind"):
            self._rebind_basic_optimizer()
        scale = float(self._inv_scale.item())
        for w, region, s in self._generic_bind:
            g = self.grad_shard[region.lo:region.hi][s]
            w.grad = g.to(device=w.device, dtype=torch.float32) * scale
        self.optimizer.step()
        for w, region, s in self._generic_bind:
            w.grad = None
            if self.dtype != torch.float32 or w.device != self.lp_shard.device:
                self.lp_shard[region.lo:region.hi][s].copy_(w.data, non_blocking=True)

    def _optimizer_step(self):
        kind = self._opt_kind
        if not self._is_fused_kind(kind):
            return self._generic_step()
        if getattr(self, "_rs_adam_ok", False):
            # units whose update already ran inside backward are done; the rest (no gradient this step) take the plain kernel
            region = self.regions[0]
            self._ensure_state(region, kind)
            for u in self.units:
                if getattr(u, "adam_done", False):
                    u.adam_done = False
                    continue
                for gi, lo, hi in self._unit_runs(u):
                    s_ = slice(lo, hi)
                    self._apply_update(kind, region, s_, self.grad_shard[s_], self.lp_shard[s_], self.param_groups[gi], self._inv_scale, 1.0)
            return
        for region in self.regions:
            self._ensure_state(region, kind)
            on_host = region.device.type == "cpu" and self.cuda
            if on_host:
                self._ensure_host_staging(region)
                # most of the gradient shard already left for the host during backward (side stream, _offload_grad_async);
                # only what was not covered then -- units without gradients, accumulation boundaries -- is copied now
                torch.cuda.current_stream().wait_stream(self.offload_stream)
                for a, b in self._uncopied_ranges(region.lo, region.hi):
                    region.host_grad[a - region.lo:b - region.lo].copy_(self.grad_shard[a:b], non_blocking=True)
                inv_host = float(self._inv_scale.item())  # also synchronises the D2H copies
                grads, lp_all, inv_dev, gs = region.host_grad, region.host_lp, None, inv_host
            else:
                grads = self.grad_shard[region.lo:region.hi]
                lp_all = self.lp_shard[region.lo:region.hi]
                if lp_all.device != region.device:  # device-side update with host-resident shards: stage the 16-bit output
                    lp_all = torch.empty(region.hi - region.lo, dtype=self.dtype, device=region.device)
                inv_dev, gs = (self._inv_scale if self.cuda else None), (1.0 if self.cuda else float(self._inv_scale.item()))
            if getattr(region, "is_nvme", False):
                self._nvme_region_step(kind, region, grads, lp_all, inv_dev, gs)
                if on_host:
                    self.lp_shard[region.lo:region.hi].copy_(region.host_lp, non_blocking=True)
                continue
            for gi, lo, hi in self.runs:
                a, b = max(lo, region.lo), min(hi, region.hi)
                if a >= b or gi < 0:
                    continue
                s = slice(a - region.lo, b - region.lo)
                hp = self.param_groups[gi]
                lp = lp_all[s] if self.dtype != torch.float32 else None
                self._apply_update(kind, region, s, grads[s], lp, hp, inv_dev, gs)
                if lp is None and on_host:
                    lp_all[s].copy_(region.master[s])
            if on_host:
                self.lp_shard[region.lo:region.hi].copy_(region.host_lp, non_blocking=True)
            elif lp_all.data_ptr() != self.lp_shard[region.lo:region.hi].data_ptr() and self.dtype != torch.float32:
                self.lp_shard[region.lo:region.hi].copy_(lp_all)

    # ---- ZeRO-Offload: gradients of host-resident optimizer regions leave the device on a side stream during backward
    def _ensure_host_staging(self, region):
        if region.host_grad is None:
            region.host_grad = torch.empty(region.hi - region.lo, dtype=self.grad_shard.dtype).pin_memory()
            region.host_lp = torch.empty(region.hi - region.lo, dtype=self.dtype).pin_memory()

    def _offload_grad_async(self, u, producer_stream):
        """Called when unit `u`'s slice of the gradient shard is final for this step (after its reduction on
        `producer_stream`): start the device -> pinned-host copy now, under the rest of backward, instead of at step()
        (reference: stage_1_and_2.py:1149-1242 overlaps its gradient offload the same way)."""
        if not self.cuda or not getattr(self, "offload_optimizer", False) or self.micro_step_id != self.gas - 1:
            return
        lo, hi = u.shard_off, u.shard_off + u.shard_numel
        stream = self.offload_stream
        started = False
        for region in self.regions:
            if region.device.type != "cpu":
                continue
            a, b = max(lo, region.lo), min(hi, region.hi)
            if a >= b:
                continue
            self._ensure_host_staging(region)
            if not started:
                stream.wait_stream(producer_stream)
                started = True
            with torch.cuda.stream(stream):
                region.host_grad[a - region.lo:b - region.lo].copy_(self.grad_shard[a:b], non_blocking=True)
            self._d2h_done.append((a, b))

    def _uncopied_ranges(self, lo, hi):
        """[lo, hi) minus everything _offload_grad_async already copied this step; clears the record."""
        done = sorted(r for r in self._d2h_done if r[1] > lo and r[0] < hi)
        self._d2h_done = [r for r in self._d2h_done if not (r[1] > lo and r[0] < hi)]
        out, cur = [], lo
        for a, b in done:
            if a > cur:
                out.append((cur, a))
            cur = max(cur, b)
        if cur < hi:
            out.append((cur, hi))
        return out

    def _nvme_region_step(self, kind, region, grads, lp_all, inv_dev, gs):
        """Sub-group by sub-group update of a file-backed region with double-buffered reads / writes."""
        from types import SimpleNamespace
        region.swap_out_all()  # no-op unless a slow path made the region resident since the last step
        sw = region.swapper
        sw.prefetch(0)
        for g, (ga, gb) in enumerate(region.bounds):
            bufs = sw.swap_in(g)
            if g + 1 < len(region.bounds):
                sw.handle.wait()  # the buffer group g+1 reads into was last written out by group g-1
                sw.prefetch(g + 1)
            view = SimpleNamespace(master=bufs["master"], state={k: v for k, v in bufs.items() if k != "master"})
            glo, ghi = region.lo + ga, region.lo + gb
            for gi, lo, hi in self.runs:
                a, b = max(lo, glo), min(hi, ghi)
                if a >= b or gi < 0:
                    continue
                s_grp = slice(a - glo, b - glo)        # inside the sub-group buffers
                s_reg = slice(a - region.lo, b - region.lo)  # inside the region-wide gradient / 16-bit buffers
                lp = lp_all[s_reg] if self.dtype != torch.float32 else None
                self._apply_update(kind, view, s_grp, grads[s_reg], lp, self.param_groups[gi], inv_dev, gs)
                if lp is None:
                    lp_all[s_reg].copy_(view.master[s_grp])
            sw.swap_out(g, async_op=True)
        sw.handle.wait()

    def _apply_update(self, kind, region, s, g, lp, hp, inv_dev, gs):
        p = region.master[s]
        st = region.state
        if kind in _ADAM_TYPES:
            if kind == "Adam":
                adamw = False
            elif kind == "AdamW":
                adamw = True
            else:
                adamw = bool(getattr(self.optimizer, "adam_w_mode", True))
            b1, b2 = hp.get("betas", (0.9, 0.999))
            flat_optim.adam_step(p, st["exp_avg"][s], st["exp_avg_sq"][s], g, lr=hp["lr"], beta1=b1, beta2=b2, eps=hp.get("eps", 1e-8),
                                 weight_decay=hp.get("weight_decay", 0.0), adamw_mode=bool(adamw),
                                 bias_correction=hp.get("bias_correction", True), step=self._step_count, lp=lp,
                                 inv_scale=inv_dev, grad_scale=gs)
        elif kind in ("FusedLion", "DeepSpeedCPULion", "Lion"):
            b1, b2 = hp.get("betas", (0.9, 0.99))
            flat_optim.lion_step(p, st["exp_avg"][s], g, lr=hp["lr"], beta1=b1, beta2=b2, weight_decay=hp.get("weight_decay", 0.0),
                                 lp=lp, inv_scale=inv_dev, grad_scale=gs)
        elif kind in ("Adagrad", "DeepSpeedCPUAdagrad"):
            flat_optim.adagrad_step(p, st["exp_avg_sq"][s], g, lr=hp["lr"], eps=hp.get("eps", 1e-10),
                                    weight_decay=hp.get("weight_decay", 0.0), lp=lp, inv_scale=inv_dev, grad_scale=gs)
        else:
            self._generic_update(region, s, g, lp, hp, inv_dev, gs)

    def _generic_update(self, region, s, g, lp, hp, inv_dev, gs):
        """Any torch.optim.Optimizer: run it on fp32 views of the master shard (reference behaviour for client optimizers)."""
        key = ("generic", region.lo, s.start, s.stop)
        if not hasattr(self, "_generic"):
            self._generic = {}
        if key not in self._generic:
            w = torch.nn.Parameter(region.master[s], requires_grad=True)
            opt_cls = type(self.optimizer)
            kw = {k: v for k, v in hp.items() if k != "params"}
            try:
                sub = opt_cls([w], **kw)
            except TypeError:
                sub = opt_cls([w], lr=hp["lr"])
            self._generic[key] = (w, sub)
        w, sub = self._generic[key]
        for k, v in hp.items():
            if k != "params":
                sub.param_groups[0][k] = v
        scale = (float(inv_dev.item()) if inv_dev is not None else 1.0) * gs
        w.grad = g.float() * scale
        sub.step()
        w.grad = None
        if lp is not None and lp.data_ptr() != w.data.data_ptr():
            lp.copy_(w.data)

    def _post_step(self):
        if self.dtype == torch.float32 and self.shard_world >= 1:
            # fp32 training: master is a separate fp32 copy; lp_shard was updated by the kernels (lp write)
            pass
        if self.p2p is not None:
            # peers may pull the updated shards only after every rank finished its step: device barrier, no host sync
            if self._qwz:
                self._p2p_refresh_qwz()
            self.p2p.barrier()
            self._p2p_shards_synced = True
            for u in self.units:
                if u.persistent:
                    self._p2p_gather_unit(u)
            self.ag_stream.wait_stream(torch.cuda.current_stream())
            self._p2p_next_slot = 0
            self._p2p_reduce_count = 0
            self._p2p_barrier_events = {}
            return
        if self.shard_world > 1 or self.offload_param:
            if self.keep_full:
                for u in self.units:
                    dist.all_gather_into_tensor(u.full, u.shard, group=self.shard_group)
            else:
                for u in self.units:
                    if u.persistent:
                        self._gather_unit_into(u.full, u)
            if self.cuda:
                self.ag_stream.wait_stream(torch.cuda.current_stream())
        elif self.cuda and self.ag_stream is not None:
            pass

    # ------------------------------------------------------------------ introspection used by engine / utils
    def get_global_grad_norm(self):
        return self._global_grad_norm

    @property
    def global_grad_norm(self):
        n = self._global_grad_norm
        return float(n.item()) if torch.is_tensor(n) else n

    def get_lr(self):
        return [g["lr"] for g in self.param_groups]

    def _master_slice(self, lo, hi):
        out = torch.empty(hi - lo, dtype=torch.float32, device=self.device)
        for r in self.regions:
            a, b = max(lo, r.lo), min(hi, r.hi)
            if a < b:
                out[a - lo:b - lo].copy_(r.master[a - r.lo:b - r.lo])
        return out

    def _state_slice(self, key, lo, hi):
        out = torch.zeros(hi - lo, dtype=torch.float32, device=self.device)
        for r in self.regions:
            a, b = max(lo, r.lo), min(hi, r.hi)
            if a < b and key in r.state:
                out[a - lo:b - lo].copy_(r.state[key][a - r.lo:b - r.lo])
        return out

    def _gather_param_from_flat(self, p, getter):
        """Full fp32 tensor for `p` assembled from every rank's shard of a rank-flat buffer."""
        u = p.ds_unit
        local = getter(u.shard_off, u.shard_off + u.shard_numel)
        if self.shard_world > 1:
            full = torch.empty(u.padded, dtype=torch.float32, device=self.device)
            dist.all_gather_into_tensor(full, local.contiguous(), group=self.shard_group)
        else:
            full = local
        off, n = u.param_slice(p.ds_index)
        return full[off:off + n].view(p.ds_shape).clone()

    def get_full_fp32_param(self, p):
        return self._gather_param_from_flat(p, self._master_slice)

    def get_full_optimizer_state(self, p, key):
        return self._gather_param_from_flat(p, lambda lo, hi: self._state_slice(key, lo, hi))

    def get_full_grad(self, p):
        inv = 1.0 / self.loss_scale if self.loss_scale else 1.0
        return self._gather_param_from_flat(p, lambda lo, hi: self.grad_shard[lo:hi].float()) * inv

    def _scatter_param_to_flat(self, p, value, setter):
        u = p.ds_unit
        off, n = u.param_slice(p.ds_index)
        lo, hi = u.shard_range()
        a, b = max(off, lo), min(off + n, hi)
        if a < b:
            setter(u.shard_off + a - lo, u.shard_off + b - lo, value.reshape(-1)[a - off:b - off].float())

    def set_full_fp32_param(self, p, value):
        def setter(lo, hi, v):
            for r in self.regions:
                a, b = max(lo, r.lo), min(hi, r.hi)
                if a < b:
                    r.master[a - r.lo:b - r.lo].copy_(v[a - lo:b - lo])
            self.lp_shard[lo:hi].copy_(v)
        self._scatter_param_to_flat(p, value.to(self.device), setter)
        u = p.ds_unit
        if self.p2p is not None:
            self._p2p_shards_written()  # on every rank, whether or not it owns a piece of p
        if self.keep_full and self.shard_world > 1:
            dist.all_gather_into_tensor(u.full, u.shard, group=self.shard_group)
        elif u.persistent and (self.shard_world > 1 or self.offload_param):
            self._gather_unit_into(u.full, u)

    # ---- local (this rank's shard) views of one parameter: flat fp32 fragments, possibly empty
    # (reference: stage3.py get_local_fp32_param / get_local_fp32_grad_for_param / set_local_hp_param)
    def _local_span(self, p):
        u = p.ds_unit
        off, n = u.param_slice(p.ds_index)
        lo, hi = u.shard_range()
        a, b = max(off, lo), min(off + n, hi)
        if a >= b:
            return None
        return u.shard_off + a - lo, u.shard_off + b - lo

    def get_local_fp32_param(self, p, optim_state_key=None):
        span = self._local_span(p)
        if span is None:
            return torch.empty(0, dtype=torch.float32, device=self.device)
        if optim_state_key is None:
            return self._master_slice(*span).clone()
        return self._state_slice(optim_state_key, *span).clone()

    def get_local_fp32_grad_for_param(self, p):
        span = self._local_span(p)
        if span is None:
            return torch.empty(0, dtype=torch.float32, device=self.device)
        inv = 1.0 / self.loss_scale if self.loss_scale else 1.0
        return self.grad_shard[span[0]:span[1]].float() * inv

    def set_local_hp_param(self, value, p, optim_state_key=None):
        span = self._local_span(p)
        if span is None:
            return
        lo, hi = span
        v = value.reshape(-1).to(self.device, torch.float32)
        assert v.numel() == hi - lo, f"local fragment has {hi - lo} elements, got {v.numel()}"
        for r in self.regions:
            a, b = max(lo, r.lo), min(hi, r.hi)
            if a >= b:
                continue
            if optim_state_key is None:
                r.master[a - r.lo:b - r.lo].copy_(v[a - lo:b - lo])
            else:
                self._ensure_state(r, self._opt_kind)
                r.state[optim_state_key][a - r.lo:b - r.lo].copy_(v[a - lo:b - lo])
        if optim_state_key is None:
            self.lp_shard[lo:hi].copy_(v)

    # names of the reference optimizers for the same operations
    def get_full_hp_param(self, p, optim_state_key=None):
        return self.get_full_fp32_param(p) if optim_state_key is None else self.get_full_optimizer_state(p, optim_state_key)

    def set_full_hp_param(self, value, p, optim_state_key=None):
        if optim_state_key is None:
            return self.set_full_fp32_param(p, value)
        return self.set_full_optimizer_state(p, value, optim_state_key)

    def get_fp32_grad_for_param(self, p):
        return self.get_full_grad(p)

    def set_lr(self, lr):
        for g in self.param_groups:
            g["lr"] = lr

    def override_loss_scale(self, loss_scale):
        """Freeze the loss scale at `loss_scale` (reference: stage3.py override_loss_scale / stage_1_and_2.py)."""
        if loss_scale != self.external_loss_scale:
            log_dist(f"[deepspeed] setting loss scale from {self.external_loss_scale} -> {loss_scale}", ranks=[0])
        self.custom_loss_scaler = True
        self.external_loss_scale = loss_scale

    def has_overflow(self, partition_gradients=True):
        return self._check_overflow()

    def check_overflow(self, partition_gradients=True):
        self.overflow = self._check_overflow()
        return self.overflow

    def scaled_global_norm(self, norm_type=2):
        assert norm_type == 2, "only L2 norm supported as of now"
        return self._compute_global_norm()

    def refresh_fp32_params(self):
        """fp32 masters <- current 16-bit parameters (after the module weights were overwritten from outside)."""
        return self.refresh_from_params()

    def initialize_optimizer_states(self):
        """Materialise the optimizer state (zeros) before the first step, e.g. for checkpoint loading into a fresh engine."""
        for r in self.regions:
            self._ensure_state(r, self._opt_kind)

    def get_lean_optimizer_state(self):
        """Per-region optimizer state without padding (reference name: _get_lean_tensors / get_lean_optimizer_state)."""
        return [{k: v for k, v in (r.state or {}).items()} for r in self.regions]

    def destroy(self):
        for h in getattr(self, "_hook_handles", []):
            h.remove()
        self._hook_handles = []
        if getattr(self, "p2p", None) is not None:
            try:
                self.p2p.close()
            except Exception:
                pass
            self.p2p = None

    def set_full_optimizer_state(self, p, value, key):
        def setter(lo, hi, v):
            for r in self.regions:
                self._ensure_state(r, self._opt_kind)
                a, b = max(lo, r.lo), min(hi, r.hi)
                if a < b:
                    r.state[key][a - r.lo:b - r.lo].copy_(v[a - lo:b - lo])
        self._scatter_param_to_flat(p, value.to(self.device), setter)

    # ------------------------------------------------------------------ checkpoint
    def _layout(self):
        return {"units": [{"name": u.name, "padded": u.padded, "shard_numel": u.shard_numel, "shard_off": u.shard_off,
                           "params": [(getattr(p, "ds_name", str(p.ds_id)), tuple(p.ds_shape), u.offsets[i])
                                      for i, p in enumerate(u.params)]} for u in self.units],
                "shard_world": self.shard_world, "total_shard": self.total_shard, "align": 64}

    def state_dict(self):
        """Per-rank optimizer shard in the reference's key layout (stage3.py:2382-2411, stage_1_and_2.py:2102-2142)
        plus `ds_unit_layout`, which tells zero_to_fp32 how this rank's flat shard maps to parameters."""
        master = self._master_slice(0, self.total_shard).cpu()
        state = {}
        keys = set()
        for r in self.regions:
            keys.update(r.state.keys())
        for k in sorted(keys):
            state[k] = self._state_slice(k, 0, self.total_shard).cpu()
        sd = {
            "zero_stage": self.stage,
            "loss_scaler": self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "partition_count": self.partition_count,
            "ds_unit_layout": self._layout(),
            "step_count": self._step_count,
            "optimizer_state_dict": {"state": {0: state},
                                     "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]},
        }
        if hasattr(self, "_generic_bind"):
            sd["basic_optimizer_state_dict"] = self.optimizer.state_dict()  # client / non-fused optimizer on the flat views
        if self.stage == 3:
            sd["fp32_flat_groups"] = [master]
        else:
            sd["single_partition_of_fp32_groups"] = [master]
            sd["base_optimizer_state"] = sd["optimizer_state_dict"]["state"]
            sd["group_paddings"] = [0]
            sd["clip_grad"] = self.clip_grad
        return sd

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=True, checkpoint_folder=None,
                        load_serial=None):
        sd = state_dict_list[self.shard_rank] if isinstance(state_dict_list, (list, tuple)) else state_dict_list
        assert sd["partition_count"] == self.partition_count, (
            f"checkpoint was saved with shard world {sd['partition_count']}, current is {self.partition_count}; "
            "use the universal checkpoint path (ds_to_universal) to reshape")
        self.loss_scaler = sd.get("loss_scaler", self.loss_scaler)
        self.dynamic_loss_scale = sd.get("dynamic_loss_scale", self.dynamic_loss_scale)
        self.overflow = sd.get("overflow", False)
        self._step_count = sd.get("step_count", 0)
        master = (sd.get("fp32_flat_groups") or sd.get("single_partition_of_fp32_groups"))[0]
        if load_from_fp32_weights:
            for r in self.regions:
                r.master.copy_(master[r.lo:r.hi])
            self.lp_shard.copy_(master.to(self.device).to(self.dtype))
        else:
            for r in self.regions:
                r.master.copy_(self.lp_shard[r.lo:r.hi].float())
        if load_optimizer_states and "basic_optimizer_state_dict" in sd and not self._is_fused_kind(self._opt_kind):
            if not hasattr(self, "_generic_bind"):
                self._rebind_basic_optimizer()
            self.optimizer.load_state_dict(sd["basic_optimizer_state_dict"])
        if load_optimizer_states:
            st = sd["optimizer_state_dict"]["state"].get(0, {})
            for r in self.regions:
                self._ensure_state(r, self._opt_kind)
                for k, v in st.items():
                    if k in r.state:
                        r.state[k].copy_(v[r.lo:r.hi])
            for g, saved in zip(self.param_groups, sd["optimizer_state_dict"]["param_groups"]):
                for k, v in saved.items():
                    g[k] = v
        self._post_load_sync()

    def _post_load_sync(self):
        if self.p2p is not None:
            torch.cuda.synchronize()
            dist.barrier(group=self.shard_group)
        if self.shard_world > 1 or self.offload_param:
            for u in self.units:
                if self.keep_full or u.persistent:
                    self._gather_unit_into(u.full, u)

    def refresh_from_params(self):
        """Re-shard after module weights were loaded into the (gathered) parameters."""
        for u in self.units:
            if u.full is not None:
                self._write_back_unit(u)

    # ------------------------------------------------------------------ checkpoint-time parameter materialisation
    def checkpoint_event_prologue(self):
        pass

    def checkpoint_event_epilogue(self):
        pass

    def gathered_unit(self, u):
        return _GatheredUnit(self, u)


class _GatheredUnit:

    def __init__(self, opt, u):
        self.opt, self.u = opt, u

    def __enter__(self):
        self.was = self.u.status
        self.opt._fetch(self.u)
        return self.u

    def __exit__(self, *exc):
        if self.was == ZeroParamStatus.NOT_AVAILABLE:
            self.opt.release_unit(self.u)
        return False


class _nullctx:

    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False
