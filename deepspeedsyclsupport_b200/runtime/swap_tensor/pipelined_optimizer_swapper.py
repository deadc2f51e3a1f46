"""Reference module name for the optimizer-state swapper (this tree has one implementation with double-buffered, i.e.
pipelined, reads and writes: optimizer_swapper.OptimizerStateSwapper)."""
from .optimizer_swapper import OptimizerStateSwapper

PartitionedOptimizerSwapper = OptimizerStateSwapper
PipelinedOptimizerSwapper = OptimizerStateSwapper
