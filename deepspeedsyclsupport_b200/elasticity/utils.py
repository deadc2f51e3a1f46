def is_torch_elastic_compatible():
    """torch.distributed.elastic exists in every torch this framework supports."""
    try:
        import torch.distributed.elastic  # noqa: F401
        return True
    except ImportError:
        return False
