"""Random-LTD schedule: how many tokens the wrapped layers keep at a given step, plus the running count of
"layer-tokens" consumed (the quantity the LR schedule can be tied to).

kept(t) = min(max_value, floor_to_multiple(min + (max - min) * t / require_steps, seq_per_step))    (fixed_linear)
consumed += global_batch * (kept * ltd_layers + max_value * (total_layers - ltd_layers))            (once per step)
State keys and config keys follow the reference (data_routing/scheduler.py) because they are checkpointed."""
import math

from ....utils.logging import logger
from .. import constants as K


class BaseScheduler(object):

    def __init__(self):
        self.state = {}

    def get_value(self, global_steps):
        kind = self.state[K.RANDOM_LTD_SCHEDULER_TYPE]
        if kind != "fixed_linear":
            raise RuntimeError("Unsupported random LTD schedule type")
        cfg = self.state[K.RANDOM_LTD_SCHEDULE_CONFIG]
        lo, hi = self.state[K.RANDOM_LTD_MIN_VALUE], self.state[K.RANDOM_LTD_MAX_VALUE]
        progress = float(global_steps) / cfg[K.RANDOM_LTD_REQUIRE_STEP]
        value = math.floor(lo + progress * (hi - lo))
        value -= value % cfg[K.RANDOM_LTD_INCREASE_STEP]
        return min(value, hi)


class RandomLTDScheduler(BaseScheduler):
    _CHECKPOINTED = (K.RANDOM_LTD_CONSUMED_LAYER_TOKENS, K.RANDOM_LTD_CURR_STEP, K.RANDOM_LTD_CURRENT_VALUE, K.RANDOM_LTD_MIN_VALUE,
                     K.RANDOM_LTD_MAX_VALUE)

    def __init__(self, config):
        super().__init__()
        self.model_layer_num = config[K.RANDOM_LTD_TOTAL_LAYER_NUM]
        self.random_ltd_layer_num = config[K.RANDOM_LTD_LAYER_NUM]
        self.config_schedule = config[K.RANDOM_LTD_SCHEDULER]
        self.global_batch_size = config[K.RANDOM_LTD_GLOBAL_BATCH_SIZE]
        lr_cfg = config.get(K.RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE) or {}
        if lr_cfg.get(K.RANDOM_LTD_LAYER_TOKEN_LR_ENABLED, False):
            logger.warning("layer-token LR schedule is not implemented")
            raise NotImplementedError
        self.reset_to_init()

    def reset_to_init(self):
        sched = self.config_schedule
        if sched is not None:
            self.state.update({
                K.RANDOM_LTD_MIN_VALUE: sched[K.RANDOM_LTD_MIN_VALUE],
                K.RANDOM_LTD_MAX_VALUE: sched[K.RANDOM_LTD_MAX_VALUE],
                K.RANDOM_LTD_CURRENT_VALUE: sched[K.RANDOM_LTD_MIN_VALUE],
                K.RANDOM_LTD_SCHEDULE_CONFIG: sched[K.RANDOM_LTD_SCHEDULE_CONFIG],
                K.RANDOM_LTD_SCHEDULER_TYPE: sched[K.RANDOM_LTD_SCHEDULER_TYPE],
            })
        self.state[K.RANDOM_LTD_CONSUMED_LAYER_TOKENS] = 0
        self.state[K.RANDOM_LTD_CURR_STEP] = -1

    def get_current_seq(self):
        return self.state[K.RANDOM_LTD_CURRENT_VALUE]

    def set_current_seq(self, seq_length):
        self.state[K.RANDOM_LTD_CURRENT_VALUE] = seq_length

    def get_random_ltd_layer_num(self):
        return self.random_ltd_layer_num

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def update_seq(self, global_steps):
        st = self.state
        if st[K.RANDOM_LTD_CURRENT_VALUE] < st[K.RANDOM_LTD_MAX_VALUE]:
            st[K.RANDOM_LTD_CURRENT_VALUE] = self.get_value(global_steps)
        if st[K.RANDOM_LTD_CURR_STEP] != global_steps:
            dropped_layers = self.random_ltd_layer_num
            full_layers = self.model_layer_num - dropped_layers
            st[K.RANDOM_LTD_CONSUMED_LAYER_TOKENS] += self.global_batch_size * (st[K.RANDOM_LTD_CURRENT_VALUE] * dropped_layers +
                                                                                st[K.RANDOM_LTD_MAX_VALUE] * full_layers)
            st[K.RANDOM_LTD_CURR_STEP] = global_steps

    def get_total_layer_tokens(self, train_iters):
        for step in range(train_iters):
            self.update_seq(step)
        return self.state[K.RANDOM_LTD_CONSUMED_LAYER_TOKENS]

    def state_dict(self):
        return {k: self.state[k] for k in self._CHECKPOINTED}

    def load_state_dict(self, state_dict):
        self.state.update({k: state_dict[k] for k in self._CHECKPOINTED})
