from ..runtime.config import CSVConfig, DeepSpeedMonitorConfig, TensorBoardConfig, WandbConfig  # noqa: F401


def get_monitor_config(param_dict):
    return DeepSpeedMonitorConfig(**{k: param_dict[k] for k in ("tensorboard", "wandb", "csv_monitor") if k in param_dict})
