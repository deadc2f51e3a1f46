"""MpiBackend (parity: deepspeed/runtime/comm/mpi.py).  The torch.distributed process group carries the same
compressed all-reduce, so the MPI flavour is the same collective on whatever backend the job was initialised with."""
from .compressed import CompressedBackend


class MpiBackend(CompressedBackend):

    def __init__(self, cuda_aware=False, mpu=None):
        super().__init__(mpu)
        self.cuda_aware = cuda_aware
