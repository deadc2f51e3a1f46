"""Exhaustive and random search over the experiment list (parity: tuner/index_based_tuner.py).  Both tuners hand out the remaining
experiments in batches; they differ only in which ones leave the pool first."""
import random

from .base_tuner import BaseTuner


class _PoolTuner(BaseTuner):

    def __init__(self, exps: list, resource_manager, metric):
        super().__init__(exps, resource_manager, metric)

    def _positions(self, count):
        """Indices into self.all_exps of the next `count` experiments."""
        raise NotImplementedError

    def next_batch(self, sample_size=1):
        take = sorted(self._positions(min(sample_size, len(self.all_exps))))
        chosen = [self.all_exps[i] for i in take]
        for i in reversed(take):
            del self.all_exps[i]
        return chosen


class RandomTuner(_PoolTuner):
    """Explore the search space in random order"""

    def _positions(self, count):
        return random.sample(range(len(self.all_exps)), count)


class GridSearchTuner(_PoolTuner):
    """Explore the search space in sequential order"""

    def _positions(self, count):
        return list(range(count))
