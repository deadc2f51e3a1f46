from .autotuner import Autotuner
