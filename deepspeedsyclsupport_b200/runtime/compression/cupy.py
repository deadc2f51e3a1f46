"""Sign-bit packing helper of the 1-bit optimizers (parity: deepspeed/runtime/compression/cupy.py:11-26).

The reference goes through CuPy for `packbits` because torch has none; here packing is part of the fused compress kernel
(csrc/quant/onebit.cu, runtime/comm/compressed.py), so this class only keeps the reference's helper surface for code that calls
it directly.  "cupy tensors" are torch tensors: both conversions are the identity and packing is done with torch ops (8 signs
per byte, most significant bit first, as numpy / cupy `packbits` lay them out)."""
import torch

_WEIGHTS = (128, 64, 32, 16, 8, 4, 2, 1)


class CupyBackend(object):

    def __init__(self):
        pass

    def torch2cupy(self, tensor):
        return tensor

    def cupy2torch(self, cupy_tensor):
        return cupy_tensor

    def packbits(self, bool_tensor):
        flat = bool_tensor.reshape(-1).to(torch.uint8)
        pad = (-flat.numel()) % 8
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        w = torch.tensor(_WEIGHTS, dtype=torch.uint8, device=flat.device)
        return (flat.view(-1, 8) * w).sum(1).to(torch.uint8)

    def unpackbits(self, packed):
        w = torch.tensor(_WEIGHTS, dtype=torch.uint8, device=packed.device)
        return ((packed.reshape(-1, 1) & w) != 0).reshape(-1)

    def compress_by_chunk(self, cupy_bool_tensor, num_chunks):
        """Pack a boolean tensor and split the bytes into `num_chunks` equal parts (one per destination rank)."""
        packed = self.packbits(cupy_bool_tensor)
        assert packed.numel() % num_chunks == 0, "packed size must divide evenly over the chunks"
        return list(torch.chunk(packed, num_chunks))
