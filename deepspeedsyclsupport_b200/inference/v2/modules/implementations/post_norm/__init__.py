from .cuda_post_ln import DSPostLNCUDAModule
