"""Exhaustive and random search over the experiment list (parity: tuner/index_based_tuner.py)."""
import random

from .base_tuner import BaseTuner


class RandomTuner(BaseTuner):
    """Explore the search space in random order"""

    def __init__(self, exps: list, resource_manager, metric):
        super().__init__(exps, resource_manager, metric)

    def next_batch(self, sample_size=1):
        sample_size = min(sample_size, len(self.all_exps))
        picked = random.sample(self.all_exps, sample_size)
        self.all_exps = [x for x in self.all_exps if x not in picked]
        return picked


class GridSearchTuner(BaseTuner):
    """Explore the search space in sequential order"""

    def __init__(self, exps: list, resource_manager, metric):
        super().__init__(exps, resource_manager, metric)

    def next_batch(self, sample_size=1):
        sample_size = min(sample_size, len(self.all_exps))
        picked, self.all_exps = self.all_exps[:sample_size], self.all_exps[sample_size:]
        return picked
