"""`MultiTensorApply` (parity: deepspeed/ops/{adam,lion}/multi_tensor_apply.py).  The reference chunks lists of tensors into
kernel launches of `chunk_size` elements; the fused optimizers here run over one flat buffer per dtype instead, so the functor is
called once per list group and the chunk size only bounds the flat views handed to it."""


class MultiTensorApply(object):

    def __init__(self, chunk_size):
        self.chunk_size = chunk_size

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
