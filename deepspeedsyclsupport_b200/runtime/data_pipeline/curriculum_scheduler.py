"""Curriculum difficulty schedules: fixed_discrete / fixed_root / fixed_linear / custom.
Parity: deepspeed/runtime/data_pipeline/curriculum_scheduler.py:11-158 (same config keys, same state dict keys)."""
import math

from ...utils.logging import logger
from .constants import *  # noqa: F401,F403

_REQUIRED = {
    CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE: (CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY, CURRICULUM_LEARNING_SCHEDULE_MAX_STEP),
    CURRICULUM_LEARNING_SCHEDULE_FIXED_ROOT: (CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP, CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP,
                                              CURRICULUM_LEARNING_SCHEDULE_ROOT_DEGREE),
    CURRICULUM_LEARNING_SCHEDULE_FIXED_LINEAR: (CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP, CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP),
    CURRICULUM_LEARNING_SCHEDULE_CUSTOM: (),
}


class CurriculumScheduler(object):

    def __init__(self, config):
        super().__init__()
        for key in (CURRICULUM_LEARNING_MIN_DIFFICULTY, CURRICULUM_LEARNING_MAX_DIFFICULTY, CURRICULUM_LEARNING_SCHEDULE_TYPE):
            assert key in config, f"Curriculum learning requires the config '{key}'"
        kind = config[CURRICULUM_LEARNING_SCHEDULE_TYPE]
        if kind not in _REQUIRED:
            raise RuntimeError("Unsupported curriculum schedule type")
        self.state = {
            CURRICULUM_LEARNING_MIN_DIFFICULTY: config[CURRICULUM_LEARNING_MIN_DIFFICULTY],
            CURRICULUM_LEARNING_MAX_DIFFICULTY: config[CURRICULUM_LEARNING_MAX_DIFFICULTY],
            CURRICULUM_LEARNING_CURRENT_DIFFICULTY: config[CURRICULUM_LEARNING_MIN_DIFFICULTY],
            CURRICULUM_LEARNING_SCHEDULE_TYPE: kind,
        }
        self.first_step = True
        self.custom_get_difficulty = None
        if kind != CURRICULUM_LEARNING_SCHEDULE_CUSTOM:
            sched = config[CURRICULUM_LEARNING_SCHEDULE_CONFIG]
            for key in _REQUIRED[kind]:
                assert key in sched, f"Curriculum learning with {kind} schedule requires the schedule_config '{key}'"
            if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE:
                diffs, steps = sched[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY], sched[CURRICULUM_LEARNING_SCHEDULE_MAX_STEP]
                # the last difficulty applies to every step after the last max_step
                assert len(steps) > 0 and len(diffs) == len(steps) + 1
            elif sched[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP] % 8 != 0:
                logger.warning("When using seqlen metric, the difficulty_step for curriculum learning has to be multiple of 8 (for FP16 data) or "
                               "16 (for INT8 data) to enable NVIDIA Tensor Core acceleration. Disregard this warning if this is unrelated to your "
                               "metric/hardware.")
            self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG] = sched

    def get_current_difficulty(self):
        return self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY]

    def set_current_difficulty(self, difficulty):
        self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] = difficulty

    def set_custom_get_difficulty(self, schedule_function):
        self.custom_get_difficulty = schedule_function

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def _discrete(self, global_steps):
        sched = self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG]
        for diff, max_step in zip(sched[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY], sched[CURRICULUM_LEARNING_SCHEDULE_MAX_STEP]):
            if global_steps <= max_step:
                return diff
        return sched[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY][-1]

    def _root(self, global_steps, root_degree=None):
        sched = self.state[CURRICULUM_LEARNING_SCHEDULE_CONFIG]
        root_degree = root_degree if root_degree is not None else sched[CURRICULUM_LEARNING_SCHEDULE_ROOT_DEGREE]
        lo, hi = self.state[CURRICULUM_LEARNING_MIN_DIFFICULTY], self.state[CURRICULUM_LEARNING_MAX_DIFFICULTY]
        frac = (float(global_steps) / sched[CURRICULUM_LEARNING_SCHEDULE_TOTAL_STEP])**(1.0 / root_degree)
        nxt = math.floor(frac * (hi - lo) + lo)
        nxt -= nxt % sched[CURRICULUM_LEARNING_SCHEDULE_DIFFICULTY_STEP]
        return min(nxt, hi)

    def get_difficulty(self, global_steps):
        kind = self.state[CURRICULUM_LEARNING_SCHEDULE_TYPE]
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_DISCRETE:
            return self._discrete(global_steps)
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_LINEAR:
            return self._root(global_steps, 1)
        if kind == CURRICULUM_LEARNING_SCHEDULE_FIXED_ROOT:
            return self._root(global_steps)
        if kind == CURRICULUM_LEARNING_SCHEDULE_CUSTOM:
            return self.custom_get_difficulty(global_steps)
        raise RuntimeError("Unsupported curriculum schedule type")

    def update_difficulty(self, global_steps):
        if self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] < self.state[CURRICULUM_LEARNING_MAX_DIFFICULTY]:
            self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY] = self.get_difficulty(global_steps)
        return self.state[CURRICULUM_LEARNING_CURRENT_DIFFICULTY]
