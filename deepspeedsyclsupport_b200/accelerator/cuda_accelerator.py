"""Reference module / class name of the CUDA accelerator (accelerator/cuda_accelerator.py CUDA_Accelerator): the single
B200 backend of this framework."""
from .real_accelerator import B200Accelerator

CUDA_Accelerator = B200Accelerator
