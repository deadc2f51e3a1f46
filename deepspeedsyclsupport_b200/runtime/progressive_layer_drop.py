"""Progressive layer dropping (PLD, arXiv:2010.13369): the keep probability of the transformer layers decays from 1 to
`theta` as  theta_t = (1 - theta) * exp(-gamma * t) + theta.  The engine calls update_state() once per optimizer step and
merges get_state() into the forward kwargs (`progressive_layer_drop`, `pld_theta`).
(reference: deepspeed/runtime/progressive_layer_drop.py)"""
import math

from ..utils.logging import log_dist


class ProgressiveLayerDrop(object):

    def __init__(self, theta=0.5, gamma=0.001):
        self.theta = float(theta)
        self.gamma = float(gamma)
        self.current_theta = 1.0
        log_dist(f"Enabled progressive layer dropping (theta = {self.theta})", ranks=[0])

    def keep_probability(self, global_step):
        return (1.0 - self.theta) * math.exp(-self.gamma * global_step) + self.theta

    def update_state(self, global_step):
        self.current_theta = self.keep_probability(global_step)

    def get_theta(self):
        return self.current_theta

    def get_state(self):
        return {"progressive_layer_drop": True, "pld_theta": self.current_theta}
