"""Availability checks of the optional monitoring back ends (parity: deepspeed/monitor/utils.py)."""
import importlib


def _require(module: str, hint: str):
    try:
        return importlib.import_module(module)
    except ImportError:
        print(hint)
        raise


def check_tb_availability():
    _require("tensorboard", "If you want to use tensorboard logging, please `pip install tensorboard`")


def check_wandb_availability():
    _require("wandb", "If you want to use wandb logging, please `pip install wandb` and follow the instructions at "
             "https://docs.wandb.ai/quickstart")
