from .cuda_pre_ln import DSPreLNCUDAModule
from .cuda_pre_rms import DSPreRMSCUDAModule
