"""Common driver of the experiment-search strategies.

A strategy subclass decides WHICH experiments to try next (`next_batch`) and may learn from finished ones (`update`);
this base class runs the batches through the ResourceManager, tracks the incumbent (best experiment and metric value,
'latency' minimised, everything else maximised) and stops after `early_stopping` experiments without improvement or
`n_trials` experiments in total.  (role of deepspeed/autotuning/tuner/base_tuner.py)"""
from ...utils.logging import logger
from ..utils import write_experiments


class BaseTuner(object):

    def __init__(self, exps, resource_manager, metric):
        self.all_exps = exps
        self.rm = resource_manager
        self.metric = metric or "throughput"
        self.best_exp = None
        self.best_metric_val = None
        self.best_iter = 0
        logger.info(f"total number of exps =  {len(self.all_exps)}")

    # -- strategy interface
    def has_next(self):
        return bool(self.all_exps)

    def next_batch(self, sample_size):
        raise NotImplementedError

    def update(self):
        pass

    # -- driver
    def _improves(self, value):
        if value is None:
            return False
        if self.best_metric_val is None:
            return True
        return value < self.best_metric_val if self.metric == "latency" else value > self.best_metric_val

    def _run_batch(self, batch):
        self.rm.schedule_experiments(write_experiments(batch, self.rm.exps_dir))
        self.rm.run()
        return self.rm.parse_results(self.metric)

    def tune(self, sample_size=1, n_trials=1000, early_stopping=None):
        done = 0
        try:
            while done < n_trials and self.has_next():
                batch = self.next_batch(sample_size)
                exp, value = self._run_batch(batch)
                if self._improves(value):
                    self.best_exp, self.best_metric_val, self.best_iter = exp, value, done
                done += len(batch)
                self.update()
                self.rm.clear()
                if early_stopping and done >= self.best_iter + early_stopping:
                    logger.info(f"Tuner early stopped at iteration {done}. Best iteration is {self.best_iter}. Early stopping threshold is {early_stopping}")
                    break
        except Exception as e:  # a broken experiment must not lose the results gathered so far
            logger.info(f"Tuner Error: {type(e).__name__}: {e}")
        return done
