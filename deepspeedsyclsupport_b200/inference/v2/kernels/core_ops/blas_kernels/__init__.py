from .blas_linear import BlasLibLinear  # noqa: F401
