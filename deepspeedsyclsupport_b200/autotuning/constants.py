"""Autotuning constants: config keys / defaults (JSON schema parity with deepspeed/autotuning/constants.py) and the
default per-stage tuning spaces."""
import os

AUTOTUNING = "autotuning"
AUTOTUNING_ENABLED = "enabled"
AUTOTUNING_FAST = "fast"
AUTOTUNING_RESULTS_DIR = "results_dir"
AUTOTUNING_RESULTS_DIR_DEFAULT = "autotuning_results"
AUTOTUNING_EXPS_DIR = "exps_dir"
AUTOTUNING_EXPS_DIR_DEFAULT = "autotuning_exps"
AUTOTUNING_OVERWRITE = "overwrite"
AUTOTUNING_START_PROFILE_STEP = "start_profile_step"
AUTOTUNING_END_PROFILE_STEP = "end_profile_step"
AUTOTUNING_METRIC_PATH = "metric_path"
AUTOTUNING_MODEL_INFO_PATH = "model_info_path"
AUTOTUNING_TUNER_TYPE = "tuner_type"
AUTOTUNING_TUNER_GRIDSEARCH = "gridsearch"
AUTOTUNING_TUNER_RANDOM = "random"
AUTOTUNING_TUNER_MODELBASED = "model_based"
AUTOTUNING_TUNER_EARLY_STOPPING = "tuner_early_stopping"
AUTOTUNING_TUNER_NUM_TRIALS = "tuner_num_trials"
AUTOTUNING_ARG_MAPPINGS = "arg_mappings"
AUTOTUNING_MAX_TRAIN_BATCH_SIZE = "max_train_batch_size"
AUTOTUNING_MIN_TRAIN_BATCH_SIZE = "min_train_batch_size"
AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU = "max_train_micro_batch_size_per_gpu"
AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU = "min_train_micro_batch_size_per_gpu"
AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES = "num_tuning_micro_batch_sizes"
AUTOTUNING_MP_SIZE = "mp_size"
AUTOTUNING_METRIC = "metric"
AUTOTUNING_METRIC_LATENCY = "latency"
AUTOTUNING_METRIC_THROUGHPUT = "throughput"
AUTOTUNING_METRIC_FLOPS = "flops"
AUTOTUNING_METRIC_FORWARD = "forward"
AUTOTUNING_METRIC_BACKWRAD = "flops"
AUTOTUNING_METRIC_STEPS = "step"
AUTOTUNING_METRIC_DEFAULT = AUTOTUNING_METRIC_THROUGHPUT
MODEL_INFO = "model_info"
MODEL_INFO_PROFILE = "profile"
MODEL_INFO_NUM_PARAMS = "num_params"
MODEL_INFO_HIDDEN_SIZE = "hidden_size"
MODEL_INFO_NUM_LAYERS = "num_layers"

DS_CONFIG = "ds_config"
BUFSIZE = 1
METRIC_PERCENT_DIFF_CONST = 0.05  # a candidate must beat the incumbent by 5% to count as an improvement
TUNING_MICRO_BATCH_SIZE_PREFIX = "z"
DEFAULT_HF_CONFIG = {"train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto", "gradient_accumulation_steps": "auto"}
DEFAULT_MIN_MEM_CONFIG = {
    "train_micro_batch_size_per_gpu": 1,
    "zero_optimization": {"stage": 3},
    "memory_break_down": False,
}

# Per-stage tuning spaces: a list value means "try each".  Sizes are in elements.  B200 note: with NVSwitch the
# reduce/all-gather bucket knobs matter far less than on PCIe/IB systems, so the spaces are small.
DEFAULT_TUNING_SPACE_ZERO_0 = {"zero_optimization": {"stage": 0}}
DEFAULT_TUNING_SPACE_ZERO_1 = {"zero_optimization": {"stage": 1, "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_bucket_size": [5e7, 5e8, 1e9]}}
DEFAULT_TUNING_SPACE_ZERO_2 = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False], "reduce_scatter": [False, True],
                                                      "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_bucket_size": [5e7, 5e8, 1e9],
                                                      "contiguous_gradients": [False, True]}}
DEFAULT_TUNING_SPACE_ZERO_3 = {"zero_optimization": {"stage": 3, "overlap_comm": [True, False], "reduce_scatter": [False, True],
                                                      "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_partitions": [True, False],
                                                      "allgather_bucket_size": [5e7, 5e8, 1e9], "contiguous_gradients": [False, True]}}
DEFAULT_TUNING_SPACES = {0: DEFAULT_TUNING_SPACE_ZERO_0, 1: DEFAULT_TUNING_SPACE_ZERO_1, 2: DEFAULT_TUNING_SPACE_ZERO_2, 3: DEFAULT_TUNING_SPACE_ZERO_3}

_TEMPLATE_DIR = os.path.join(os.path.dirname(os.path.realpath(__file__)), "config_templates")
DEFAULT_TEMPLATE_PATH_ZERO_0 = os.path.join(_TEMPLATE_DIR, "template_zero0.json")
DEFAULT_TEMPLATE_PATH_ZERO_1 = os.path.join(_TEMPLATE_DIR, "template_zero1.json")
DEFAULT_TEMPLATE_PATH_ZERO_2 = os.path.join(_TEMPLATE_DIR, "template_zero2.json")
DEFAULT_TEMPLATE_PATH_ZERO_3 = os.path.join(_TEMPLATE_DIR, "template_zero3.json")
