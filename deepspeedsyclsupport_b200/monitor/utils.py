"""(parity: deepspeed/monitor/utils.py)"""


def check_tb_availability():
    try:
        import tensorboard  # noqa: F401
    except ImportError:
        print('If you want to use tensorboard logging, please `pip install tensorboard`')
        raise


def check_wandb_availability():
    try:
        import wandb  # noqa: F401
    except ImportError:
        print('If you want to use wandb logging, please `pip install wandb` and follow the instructions at https://docs.wandb.ai/quickstart')
        raise
