"""LR schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR, WarmupCosineLR.
Parity: deepspeed/runtime/lr_schedules.py:267,370,634,723,774 (same constructor arguments and curves)."""
import argparse
import math

from torch.optim import Optimizer

from ..utils.logging import logger

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST = "LRRangeTest"
ONE_CYCLE = "OneCycle"
WARMUP_LR = "WarmupLR"
WARMUP_DECAY_LR = "WarmupDecayLR"
WARMUP_COSINE_LR = "WarmupCosineLR"
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR, WARMUP_COSINE_LR]

WARMUP_LOG_RATE = "log"
WARMUP_LINEAR_RATE = "linear"


def add_tuning_arguments(parser):
    group = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    group.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    group.add_argument("--lr_range_test_min_lr", type=float, default=0.001)
    group.add_argument("--lr_range_test_step_rate", type=float, default=1.0)
    group.add_argument("--lr_range_test_step_size", type=int, default=1000)
    group.add_argument("--lr_range_test_staircase", type=bool, default=False)
    group.add_argument("--cycle_first_step_size", type=int, default=1000)
    group.add_argument("--cycle_first_stair_count", type=int, default=-1)
    group.add_argument("--cycle_second_step_size", type=int, default=-1)
    group.add_argument("--cycle_second_stair_count", type=int, default=-1)
    group.add_argument("--decay_step_size", type=int, default=1000)
    group.add_argument("--cycle_min_lr", type=float, default=0.01)
    group.add_argument("--cycle_max_lr", type=float, default=0.1)
    group.add_argument("--decay_lr_rate", type=float, default=0.0)
    group.add_argument("--cycle_momentum", default=False, action="store_true")
    group.add_argument("--cycle_min_mom", type=float, default=0.8)
    group.add_argument("--cycle_max_mom", type=float, default=0.9)
    group.add_argument("--decay_mom_rate", type=float, default=0.0)
    group.add_argument("--warmup_min_lr", type=float, default=0)
    group.add_argument("--warmup_max_lr", type=float, default=0.001)
    group.add_argument("--warmup_num_steps", type=int, default=1000)
    group.add_argument("--warmup_type", type=str, default=WARMUP_LOG_RATE)
    return parser


# command-line flag -> schedule parameter, per schedule family (flag names are the parameter names)
_ARG_PARAMS = {
    LR_RANGE_TEST: ("lr_range_test_min_lr", "lr_range_test_step_rate", "lr_range_test_step_size",
                    "lr_range_test_staircase"),
    ONE_CYCLE: ("cycle_first_step_size", "cycle_first_stair_count", "cycle_second_step_size",
                "cycle_second_stair_count", "decay_step_size", "cycle_min_lr", "cycle_max_lr", "decay_lr_rate",
                "cycle_min_mom", "cycle_max_mom", "decay_mom_rate"),
    "warmup": ("warmup_min_lr", "warmup_max_lr", "warmup_num_steps", "warmup_type"),
}


def parse_arguments():
    """Parse the convergence-tuning flags out of ``sys.argv`` (parity: runtime/lr_schedules.py:158-164)."""
    import argparse
    parser = add_tuning_arguments(argparse.ArgumentParser())
    lr_sched_args, unknown_args = parser.parse_known_args()
    return lr_sched_args, unknown_args


def get_config_from_args(args):
    """Build a ``{"type", "params"}`` scheduler config from parsed tuning flags; returns ``(config, error)``
    (parity: runtime/lr_schedules.py:212-232).  Flags left at ``None`` are not copied."""
    name = getattr(args, LR_SCHEDULE, None)
    if name is None:
        return None, "--{} not specified on command line".format(LR_SCHEDULE)
    if name not in VALID_LR_SCHEDULES:
        return None, "{} is not supported LR schedule".format(name)
    family = name if name in (LR_RANGE_TEST, ONE_CYCLE) else "warmup"
    params = {k: getattr(args, k) for k in _ARG_PARAMS[family] if getattr(args, k, None) is not None}
    return {"type": name, "params": params}, None


def get_lr_from_config(config):
    """The learning rate a scheduler config starts from / peaks at; returns ``(lr, error)``
    (parity: runtime/lr_schedules.py:235-255)."""
    if "type" not in config:
        return None, "LR schedule type not defined in config"
    if "params" not in config:
        return None, "LR schedule params not defined in config"
    name, params = config["type"], config["params"]
    if name not in VALID_LR_SCHEDULES:
        return None, "{} is not a valid LR schedule".format(name)
    key = {LR_RANGE_TEST: "lr_range_test_min_lr", ONE_CYCLE: "cycle_max_lr"}.get(name, "warmup_max_lr")
    return params[key], ""


def get_torch_optimizer(optimizer):
    if isinstance(optimizer, Optimizer):
        return optimizer
    if hasattr(optimizer, "optimizer") and isinstance(optimizer.optimizer, Optimizer):
        return optimizer.optimizer
    if hasattr(optimizer, "param_groups"):
        return optimizer
    raise TypeError("{} is not a subclass of torch.optim.Optimizer".format(type(optimizer).__name__))


class _Base:

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]

    def get_last_lr(self):
        assert getattr(self, "_last_lr", None) is not None, "need to call step() first"
        return self._last_lr

    def _apply(self, lrs):
        for param_group, lr in zip(self.optimizer.param_groups, lrs):
            param_group["lr"] = lr
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]


class LRRangeTest(_Base):

    def __init__(self, optimizer, lr_range_test_min_lr=1e-3, lr_range_test_step_size=2000,
                 lr_range_test_step_rate=1.0, lr_range_test_staircase=False, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        if isinstance(lr_range_test_min_lr, (list, tuple)):
            if len(lr_range_test_min_lr) != len(self.optimizer.param_groups):
                raise ValueError("Expected {} lr_range_test_min_lr, got {}".format(len(self.optimizer.param_groups), len(lr_range_test_min_lr)))
            self.min_lr = list(lr_range_test_min_lr)
        else:
            self.min_lr = [lr_range_test_min_lr] * len(self.optimizer.param_groups)
        self.step_size = lr_range_test_step_size
        self.step_rate = lr_range_test_step_rate
        self.last_batch_iteration = last_batch_iteration
        self.staircase = lr_range_test_staircase
        self.interval_fn = self._staircase_interval if lr_range_test_staircase else self._continuous_interval
        if last_batch_iteration == -1:
            self._apply(self.min_lr)

    def _staircase_interval(self):
        return math.floor(float(self.last_batch_iteration + 1) / self.step_size)

    def _continuous_interval(self):
        return float(self.last_batch_iteration + 1) / self.step_size

    def _get_increase(self):
        return 1 + self.step_rate * self.interval_fn()

    def get_lr(self):
        inc = self._get_increase()
        return [m * inc for m in self.min_lr]

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply(self.get_lr())


class OneCycle(_Base):

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate=0.0, cycle_first_step_size=2000,
                 cycle_second_step_size=None, cycle_first_stair_count=0, cycle_second_stair_count=None,
                 decay_step_size=0, cycle_momentum=True, cycle_min_mom=0.8, cycle_max_mom=0.9, decay_mom_rate=0.0,
                 last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        cycle_first_step_size = float(cycle_first_step_size)
        cycle_second_step_size = float(cycle_second_step_size) if cycle_second_step_size is not None else cycle_first_step_size
        self.total_size = cycle_first_step_size + cycle_second_step_size
        self.step_ratio = cycle_first_step_size / self.total_size
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None else cycle_second_stair_count
        self.decay_step_size = decay_step_size
        if math.isclose(self.decay_step_size, 0):
            self.skip_lr_decay = True
            self.skip_mom_decay = True
        else:
            self.skip_lr_decay = False
            self.skip_mom_decay = False
        self.min_lrs = [cycle_min_lr] * len(self.optimizer.param_groups)
        if last_batch_iteration == -1:
            for lr, group in zip(self.min_lrs, self.optimizer.param_groups):
                group["lr"] = lr
        self.max_lrs = [cycle_max_lr] * len(self.optimizer.param_groups)
        self.decay_lr_rate = decay_lr_rate
        if math.isclose(self.decay_lr_rate, 0):
            self.skip_lr_decay = True
        self.cycle_momentum = cycle_momentum
        if cycle_momentum:
            if "betas" not in self.optimizer.defaults:
                logger.warning(f"cycle_momentum is disabled because optimizer {type(self.optimizer).__name__} does not support momentum")
                self.cycle_momentum = False
            else:
                self.decay_mom_rate = decay_mom_rate
                self.min_moms = [(cycle_min_mom, 0.99)] * len(self.optimizer.param_groups)
                self.max_moms = [(cycle_max_mom, 0.99)] * len(self.optimizer.param_groups)
                if last_batch_iteration == -1:
                    for momentum, group in zip(self.min_moms, self.optimizer.param_groups):
                        group["betas"] = momentum
                if math.isclose(self.decay_mom_rate, 0):
                    self.skip_mom_decay = True
        self.last_batch_iteration = last_batch_iteration
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def _get_scale_factor(self):
        batch_iteration = self.last_batch_iteration + 1
        cycle = math.floor(1 + batch_iteration / self.total_size)
        x = 1.0 + batch_iteration / self.total_size - cycle
        if x <= self.step_ratio:
            return x / self.step_ratio
        return (x - 1) / (self.step_ratio - 1)

    def _get_cycle_lr(self):
        sf = self._get_scale_factor()
        return [mn + (mx - mn) * sf for mn, mx in zip(self.min_lrs, self.max_lrs)]

    def _get_cycle_mom(self):
        sf = self._get_scale_factor()
        return [(mx[0] - (mx[0] - mn[0]) * sf, beta1) for (mn, mx), (_, beta1) in
                zip(zip(self.min_moms, self.max_moms), self.min_moms)]

    def _get_decay_lr(self, decay_batch_iteration):
        decay_interval = decay_batch_iteration / self.decay_step_size
        lr_decay_factor = 1 + self.decay_lr_rate * decay_interval
        return [mn / lr_decay_factor for mn in self.min_lrs]

    def _get_decay_mom(self, decay_batch_iteration):
        if self.skip_mom_decay:
            return self.max_moms
        decay_interval = decay_batch_iteration / self.decay_step_size
        mom_decay_factor = 1 + self.decay_mom_rate * decay_interval
        return [(beta0 * mom_decay_factor, beta1) for beta0, beta1 in self.max_moms]

    def get_lr(self):
        if self.last_batch_iteration < self.total_size:
            return self._get_cycle_lr()
        if self.skip_lr_decay:
            return self.min_lrs
        return self._get_decay_lr(self.last_batch_iteration - self.total_size + 1)

    def get_mom(self):
        if not self.cycle_momentum:
            return None
        if self.last_batch_iteration < self.total_size:
            return self._get_cycle_mom()
        return self._get_decay_mom(self.last_batch_iteration - self.total_size + 1)

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply(self.get_lr())
        if self.cycle_momentum:
            for param_group, momentum in zip(self.optimizer.param_groups, self.get_mom()):
                param_group["betas"] = momentum


class WarmupLR(_Base):

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.min_lrs = self._format_param(self.optimizer, warmup_min_lr, "min_lr")
        self.max_lrs = self._format_param(self.optimizer, warmup_max_lr, "max_lr")
        self.delta_lrs = [big - small for big, small in zip(self.max_lrs, self.min_lrs)]
        self.warmup_num_steps = max(2, warmup_num_steps)
        if warmup_type not in {WARMUP_LOG_RATE, WARMUP_LINEAR_RATE}:
            logger.warning(f"Using unknown warmup_type: {warmup_type}. The increasing function is set to default (log)")
            warmup_type = WARMUP_LOG_RATE
        self.warmup_type = warmup_type
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        self.last_batch_iteration = last_batch_iteration

    def get_lr(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return [0.0]
        gamma = self._get_gamma()
        return [min_lr + (delta_lr * gamma) for min_lr, delta_lr in zip(self.min_lrs, self.delta_lrs)]

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        self._apply(self.get_lr())

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            return self.last_batch_iteration / self.warmup_num_steps
        return 1.0

    def _format_param(self, optimizer, param_value, param_name):
        if isinstance(param_value, (list, tuple)):
            if len(param_value) != len(optimizer.param_groups):
                raise ValueError("expected {} value for {}, got {}".format(len(optimizer.param_groups), param_name, FileNotFoundError(param_value)))
            return list(param_value)
        return [param_value] * len(optimizer.param_groups)


class WarmupDecayLR(WarmupLR):

    def __init__(self, optimizer, total_num_steps, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, warmup_type, last_batch_iteration)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning("total_num_steps {} is less than warmup_num_steps {}".format(total_num_steps, warmup_num_steps))

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return super()._get_gamma()
        return max(0.0, float(self.total_num_steps - self.last_batch_iteration) /
                   float(max(1.0, self.total_num_steps - self.warmup_num_steps)))


class WarmupCosineLR(_Base):

    def __init__(self, optimizer, total_num_steps, warmup_min_ratio=0.0, warmup_num_steps=1000, cos_min_ratio=0.0001,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.total_num_steps = total_num_steps
        self.last_batch_iteration = last_batch_iteration
        self.cos_min_ratio = cos_min_ratio
        self.warmup_type = warmup_type if warmup_type in {WARMUP_LOG_RATE, WARMUP_LINEAR_RATE} else WARMUP_LOG_RATE
        self.warmup_min_ratio = warmup_min_ratio
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning("total_num_steps {} is less than warmup_num_steps {}".format(total_num_steps, warmup_num_steps))
        self.org_lrs = [group["lr"] for group in self.optimizer.param_groups]

    def get_lr_ratio(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return 0.0
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                ratio = self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            else:
                ratio = self.last_batch_iteration / self.warmup_num_steps
            ratio_delta = 1.0 - self.warmup_min_ratio
            return self.warmup_min_ratio + ratio * ratio_delta
        real_last_step = self.last_batch_iteration - self.warmup_num_steps + 1
        real_total_steps = self.total_num_steps - self.warmup_num_steps
        ratio_delta = 1.0 - self.cos_min_ratio
        ratio = (1 + math.cos(math.pi * real_last_step / real_total_steps)) / 2
        return max(0.0, self.cos_min_ratio + ratio_delta * ratio)

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        self._apply(self.get_lr())

    def get_lr(self):
        if self.last_batch_iteration < 0:
            return [0.0]
        r = self.get_lr_ratio()
        return [org_lr * r for org_lr in self.org_lrs]
