"""Parallel building blocks that go beyond torch.distributed: the peer-memory (symmetric memory) runtime and the
collectives / fused compute+collective kernels built on it.  Strategy-level APIs live where the reference has them
(`moe/`, `sequence/`, `runtime/pipe/`, `module_inject/`)."""
from .symm import SymmetricBuffer, EmulatedPeers  # noqa: F401
