"""compression_scheduler: switches each technique on for its modules once training reaches the technique's
schedule_offset (sparse pruning also has an end offset).  step() is called by the engine after every optimizer step.
API parity: deepspeed/compression/scheduler.py."""
from ..utils.logging import logger
from . import constants as C
from .compress import get_module_name
from .helper import recursive_getattr

# technique -> (module flag it raises, has an end offset)
_FLAGS = {
    C.WEIGHT_QUANTIZATION: ("weight_quantization_enabled", False),
    C.ACTIVATION_QUANTIZATION: ("activation_quantization_enabled", False),
    C.SPARSE_PRUNING: ("sparse_pruning_enabled", True),
    C.HEAD_PRUNING: ("head_pruning_enabled", False),
    C.ROW_PRUNING: ("row_pruning_enabled", False),
    C.CHANNEL_PRUNING: ("channel_pruning_enabled", False),
}


class compression_scheduler():

    def __init__(self, model, compression_config):
        self.model = model
        self.compression_config = compression_config
        self.make_init()
        self.training_steps = 0
        self.weight_quantization_enabled = False
        self.verbose = {t: False for t in _FLAGS}

    def make_init(self):
        self.different_compression_methods = {}
        for technique, content in self.compression_config.items():
            if C.LAYER_REDUCTION in technique:
                continue
            shared = content[C.SHARED_PARAMETERS]
            entry = {C.TECHNIQUE_ENABLED: shared[C.TECHNIQUE_ENABLED], C.SHARED_PARAMETERS: shared, C.DIFFERENT_GROUPS: []}
            claimed = set()
            for group_name, group in content[C.DIFFERENT_GROUPS].items():
                names = []
                for key_word in group[C.DIFFERENT_GROUPS_MODULE_SCOPE]:
                    matched, claimed = get_module_name(group_name, self.model, key_word, claimed, verbose=False)
                    names.extend(matched)
                if names:
                    entry[C.DIFFERENT_GROUPS].append([group_name, names, group[C.DIFFERENT_GROUPS_PARAMETERS]])
            self.different_compression_methods[technique] = entry

    def _check(self, technique):
        entry = self.different_compression_methods.get(technique)
        if entry is None or not entry[C.TECHNIQUE_ENABLED]:
            return
        flag, has_end = _FLAGS[technique]
        shared = entry[C.SHARED_PARAMETERS]
        if self.training_steps < shared[C.TECHNIQUE_SCHEDULE_OFFSET]:
            return
        if has_end and self.training_steps > shared[C.TECHNIQUE_SCHEDULE_OFFSET_END]:
            return
        for _, names, _ in entry[C.DIFFERENT_GROUPS]:
            for name in names:
                setattr(recursive_getattr(self.model, name), flag, True)
        if not self.verbose[technique]:
            logger.info(f"{technique.replace('_', ' ').capitalize()} is enabled at step {self.training_steps}")
            self.verbose[technique] = True
            if technique == C.WEIGHT_QUANTIZATION:
                self.weight_quantization_enabled = True

    # per-technique names kept for callers of the old API
    def check_weight_quantization(self):
        self._check(C.WEIGHT_QUANTIZATION)

    def check_activation_quantization(self):
        self._check(C.ACTIVATION_QUANTIZATION)

    def check_sparse_pruning(self):
        self._check(C.SPARSE_PRUNING)

    def check_head_pruning(self):
        self._check(C.HEAD_PRUNING)

    def check_row_pruning(self):
        self._check(C.ROW_PRUNING)

    def check_channel_pruning(self):
        self._check(C.CHANNEL_PRUNING)

    def check_all_modules(self):
        for technique in _FLAGS:
            self._check(technique)

    def step(self, step_zero_check=False):
        if not step_zero_check:
            self.training_steps += 1
        self.check_all_modules()
