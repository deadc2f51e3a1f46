"""RandomLayerTokenDrop: wraps a transformer layer so that, while training, it only processes a random subset of the
tokens (sorted, so causal order survives) and scatters its output back into the full sequence.
Parity: deepspeed/runtime/data_pipeline/data_routing/basic_layer.py:13-113; token movement runs on the csrc/random_ltd
kernels through ops/random_ltd."""
from torch import Tensor
from torch.nn import Module

from ....ops.random_ltd.dropping_utils import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens
from ....utils.logging import logger
from ..constants import *  # noqa: F401,F403


class RandomLayerTokenDrop(Module):
    """A layer wrapper for random LTD"""

    def __init__(self, layer: Module):
        super(RandomLayerTokenDrop, self).__init__()
        self.random_ltd_layer = layer
        self.random_ltd_scheduler = None
        self.max_length = None
        self.reserved_length = -1
        self.curr_seq = -1
        self.batch_first = False

    def init_config(self, config, scheduler, random_ltd_layer_id):
        self.random_ltd_scheduler = scheduler
        self.random_ltd_layer_id = random_ltd_layer_id
        self.max_length = self.random_ltd_scheduler.state[RANDOM_LTD_MAX_VALUE]
        self.mask_name = config[RANDOM_LTD_MODEL_MASK_NAME]
        self.micro_bs = config[RANDOM_LTD_MICRO_BATCH_SIZE]
        self.random_ltd_num_layer = self.random_ltd_scheduler.random_ltd_layer_num
        hs_order = config[RANDOM_LTD_HIDDEN_STATE_ORDER]
        self.model_type = config[RANDOM_LTD_MODEL_TYPE]
        if hs_order == "batch_seq_dim":
            self.batch_first = True
        elif hs_order == "seq_batch_dim":
            self.batch_first = False
        else:
            logger.warning("************For now, we only support batch_seq_dim or seq_batch_dim inputs.************")
            raise NotImplementedError
        if self.model_type == "encoder":
            self.index_generator = bert_sample_tokens
        elif self.model_type == "decoder":
            self.index_generator = gpt_sample_tokens
        else:
            logger.warning("************For now, we only support encoder-only or decoder-only models************")
            raise NotImplementedError

    def get_hidden_tensor_shape(self, hidden_states):
        if self.batch_first:
            self.curr_micro_batch, self.curr_seq = hidden_states.size()[0], hidden_states.size()[1]
        else:
            self.curr_seq, self.curr_micro_batch = hidden_states.size()[0], hidden_states.size()[1]

    def forward(self, hidden_states, **kwargs) -> Tensor:
        if self.random_ltd_scheduler is not None:
            self.reserved_length = self.random_ltd_scheduler.get_current_seq()
            self.get_hidden_tensor_shape(hidden_states)
        if not (self.training and self.random_ltd_scheduler is not None and self.reserved_length < self.curr_seq):
            return self.random_ltd_layer(hidden_states, **kwargs)
        mask = kwargs[self.mask_name] if self.mask_name is not None else None
        state = self.random_ltd_scheduler.state
        if self.random_ltd_layer_id == 0:
            # the first wrapped layer draws the indices of ALL wrapped layers for this step
            sampled_indices, part_attention_mask = self.index_generator(self.reserved_length, self.curr_seq, self.curr_micro_batch,
                                                                        self.random_ltd_num_layer, hidden_states.device, mask)
            state[RANDOM_LTD_SAMPLE_INDEX] = sampled_indices
            state[RANDOM_LTD_ATTENTION_MASK] = part_attention_mask
        else:
            sampled_indices = state[RANDOM_LTD_SAMPLE_INDEX]
            part_attention_mask = state[RANDOM_LTD_ATTENTION_MASK]
        my_indices = sampled_indices[self.random_ltd_layer_id, :, :]
        hidden_states, part_hidden_states = GatherTokens.apply(hidden_states, my_indices, self.batch_first)
        if self.mask_name is not None:
            kwargs[self.mask_name] = part_attention_mask[self.random_ltd_layer_id] if self.model_type == "encoder" else part_attention_mask
        outputs = self.random_ltd_layer(part_hidden_states, **kwargs)
        if isinstance(outputs, tuple):
            hidden_states = ScatterTokens.apply(hidden_states, outputs[0], my_indices, self.batch_first)
            return (hidden_states, ) + tuple(outputs[1:])
        if isinstance(outputs, Tensor):
            return ScatterTokens.apply(hidden_states, outputs, my_indices, self.batch_first)
        logger.warning("************For now, we only support tuple and tensor output.  You need to adjust the output according to the layer in your model************")
        raise NotImplementedError
