"""Model-based tuner: evaluate a few random configs, fit the surrogate cost model on (config features -> metric), then
repeatedly run the configs the model ranks highest (with an epsilon of random exploration) and refit.
Parity: deepspeed/autotuning/tuner/model_based_tuner.py."""
import json
import numbers
import os
import random

import numpy as np

from ..constants import AUTOTUNING, AUTOTUNING_METRIC_PATH
from ..utils import flatten
from .base_tuner import BaseTuner
from .cost_model import XGBoostCostModel

INIT_NUM = 2


class ModelBasedTuner(BaseTuner):
    """Exploring the search space with a cost model"""

    def __init__(self, exps: list, resource_manager, metric, tuning_space):
        super().__init__(exps, resource_manager, metric)
        self.tuning_space = tuning_space
        self.best_iter = 0
        self.all_configs = [e["ds_config"] for e in exps]
        self.num_all_configs = len(self.all_configs)
        self.dims = [len(self.all_configs)]
        self.visited = set()
        self.trials = []
        self.trial_pt = 0
        self.evaluated_configs = []
        self.evaluated_perf = []
        self.cost_model = XGBoostCostModel("rank")
        self._feature_keys = sorted({k for c in self.all_configs for k, v in flatten(c).items() if isinstance(v, (numbers.Number, bool))})
        init = list(range(min(INIT_NUM, self.num_all_configs)))
        random.shuffle(init)
        for idx in init:
            self.visited.add(idx)
            self.trials.append(idx)

    def _features(self, config):
        flat = flatten(config)
        return [float(flat.get(k, 0) or 0) for k in self._feature_keys]

    def find_estimated_top_configs(self):
        """Indices of all configs, best predicted first (latency is minimised)."""
        preds = self.cost_model.predict([self._features(c) for c in self.all_configs])
        order = np.argsort(preds)
        return list(order if self.metric == "latency" else order[::-1])

    def next_batch(self, sample_size):
        sampled = []
        while len(sampled) < sample_size and len(self.visited) <= self.num_all_configs:
            if self.trial_pt < len(self.trials):
                idx = self.trials[self.trial_pt]
                self.trial_pt += 1
            else:
                remaining = [i for i in range(self.num_all_configs) if i not in self.visited]
                if not remaining:
                    break
                idx = random.choice(remaining)
                self.visited.add(idx)
            sampled.append(self.all_exps_by_index(idx))
        return sampled

    def all_exps_by_index(self, idx):
        return self._exp_list[idx]

    def has_next(self):
        return len(self.evaluated_configs) + (len(self.trials) - self.trial_pt) < self.num_all_configs or self.trial_pt < len(self.trials)

    @property
    def _exp_list(self):
        if not hasattr(self, "_exps_frozen"):
            self._exps_frozen = list(self.all_exps)
        return self._exps_frozen

    def update(self):
        for exp_id, (exp, err) in self.rm.finished_experiments.items():
            idx = self.all_configs.index(exp["ds_config"]) if exp["ds_config"] in self.all_configs else None
            if idx is None:
                continue
            self.evaluated_configs.append(self._features(exp["ds_config"]))
            metric_file = exp["ds_config"][AUTOTUNING][AUTOTUNING_METRIC_PATH]
            if not err and os.path.exists(metric_file):
                with open(metric_file, "r") as f:
                    self.evaluated_perf.append(json.load(f)[self.metric])
            else:
                self.evaluated_perf.append(0.0)
        if not self.evaluated_perf:
            return
        self.cost_model.fit(self.evaluated_configs, self.evaluated_perf)
        estimated = self.find_estimated_top_configs()
        # queue the best-ranked unvisited configs; every 4th pick is random exploration
        for n, idx in enumerate(int(i) for i in estimated):
            if len(self.trials) - self.trial_pt >= 2:
                break
            if n % 4 == 3:
                rest = [i for i in range(self.num_all_configs) if i not in self.visited]
                idx = random.choice(rest) if rest else idx
            if idx not in self.visited:
                self.visited.add(idx)
                self.trials.append(idx)
