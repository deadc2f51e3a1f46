"""Progressive layer dropping schedule: theta(t) = (1 - theta) * exp(-gamma * t) + theta.
Parity: deepspeed/runtime/progressive_layer_drop.py:10-40."""
import numpy as np

from ..utils.logging import log_dist


class ProgressiveLayerDrop(object):
    r"""Progressive Layer Dropping (PLD) for model training (https://arxiv.org/pdf/2010.13369.pdf).

    Args:
        theta (float): a hyper-parameter that controls the trade-off between training time and robustness (lower = faster).
        gamma (float): a hyper-parameter that controls how fast the drop ratio increases.
    """

    def __init__(self, theta=0.5, gamma=0.001):
        super().__init__()
        self.theta = theta
        self.gamma = gamma
        self.current_theta = 1.0
        log_dist(f"Enabled progressive layer dropping (theta = {self.theta})", ranks=[0])

    def get_state(self):
        kwargs = {"progressive_layer_drop": True, "pld_theta": self.get_theta()}
        return kwargs

    def get_theta(self):
        return self.current_theta

    def update_state(self, global_step):

        def _prob(x, gamma, p):
            return (1. - p) * np.exp(-gamma * x) + p

        self.current_theta = _prob(global_step, self.gamma, self.theta)
