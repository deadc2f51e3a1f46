"""Random layer-token drop (random-LTD): a wrapper that lets a transformer layer see only a random, order-preserving subset
of the sequence while training and writes the layer's output back into the full-length hidden states.

What the wrapper does per forward (role of deepspeed/runtime/data_pipeline/data_routing/basic_layer.py:13-113):
    kept = scheduler.get_current_seq()            tokens this layer may process at the current step
    layer 0 draws sorted index sets for ALL wrapped layers (one launch), the others read theirs from the scheduler state
    gather kept tokens  ->  wrapped layer  ->  scatter results into the untouched full sequence
Token movement runs on the csrc/random_ltd kernels through ops/random_ltd (GatherTokens / ScatterTokens autograd functions).
"""
from torch import Tensor
from torch.nn import Module

from ....ops.random_ltd.dropping_utils import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens
from ....utils.logging import logger
from ..constants import *  # noqa: F401,F403

_SAMPLERS = {"encoder": bert_sample_tokens, "decoder": gpt_sample_tokens}
_LAYOUTS = {"batch_seq_dim": True, "seq_batch_dim": False}  # hidden-state order -> batch_first


def _unsupported(what):
    logger.warning(f"************For now, we only support {what}************")
    return NotImplementedError(what)


class RandomLayerTokenDrop(Module):
    """A layer wrapper for random LTD"""

    def __init__(self, layer: Module):
        super().__init__()
        self.random_ltd_layer = layer
        self.random_ltd_scheduler = None
        self.max_length = None
        self.reserved_length = self.curr_seq = -1
        self.batch_first = False

    def init_config(self, config, scheduler, random_ltd_layer_id):
        layout, family = config[RANDOM_LTD_HIDDEN_STATE_ORDER], config[RANDOM_LTD_MODEL_TYPE]
        if layout not in _LAYOUTS:
            raise _unsupported("batch_seq_dim or seq_batch_dim inputs.")
        if family not in _SAMPLERS:
            raise _unsupported("encoder-only or decoder-only models")
        self.batch_first, self.model_type, self.index_generator = _LAYOUTS[layout], family, _SAMPLERS[family]
        self.random_ltd_scheduler, self.random_ltd_layer_id = scheduler, random_ltd_layer_id
        self.random_ltd_num_layer = scheduler.random_ltd_layer_num
        self.max_length = scheduler.state[RANDOM_LTD_MAX_VALUE]
        self.mask_name = config[RANDOM_LTD_MODEL_MASK_NAME]
        self.micro_bs = config[RANDOM_LTD_MICRO_BATCH_SIZE]

    def get_hidden_tensor_shape(self, hidden_states):
        a, b = hidden_states.shape[:2]
        self.curr_micro_batch, self.curr_seq = (a, b) if self.batch_first else (b, a)

    def _indices_for_this_step(self, hidden_states, mask):
        """(index sets [layers, batch, kept], attention mask(s) for the kept tokens): drawn by wrapped layer 0, shared through
        the scheduler state with the layers behind it."""
        state = self.random_ltd_scheduler.state
        if self.random_ltd_layer_id == 0:
            state[RANDOM_LTD_SAMPLE_INDEX], state[RANDOM_LTD_ATTENTION_MASK] = self.index_generator(
                self.reserved_length, self.curr_seq, self.curr_micro_batch, self.random_ltd_num_layer, hidden_states.device, mask)
        return state[RANDOM_LTD_SAMPLE_INDEX], state[RANDOM_LTD_ATTENTION_MASK]

    def forward(self, hidden_states, **kwargs) -> Tensor:
        sched = self.random_ltd_scheduler
        if sched is not None:
            self.reserved_length = sched.get_current_seq()
            self.get_hidden_tensor_shape(hidden_states)
        dropping = self.training and sched is not None and self.reserved_length < self.curr_seq
        if not dropping:
            return self.random_ltd_layer(hidden_states, **kwargs)
        has_mask = self.mask_name is not None
        all_indices, kept_mask = self._indices_for_this_step(hidden_states, kwargs[self.mask_name] if has_mask else None)
        mine = all_indices[self.random_ltd_layer_id, :, :]
        full, kept = GatherTokens.apply(hidden_states, mine, self.batch_first)
        if has_mask:  # encoders get a mask per layer, decoders share one causal mask of the kept length
            kwargs[self.mask_name] = kept_mask[self.random_ltd_layer_id] if self.model_type == "encoder" else kept_mask
        result = self.random_ltd_layer(kept, **kwargs)
        if isinstance(result, Tensor):
            return ScatterTokens.apply(full, result, mine, self.batch_first)
        if isinstance(result, tuple):
            return (ScatterTokens.apply(full, result[0], mine, self.batch_first), *result[1:])
        raise _unsupported("tuple and tensor output.  You need to adjust the output according to the layer in your model")
