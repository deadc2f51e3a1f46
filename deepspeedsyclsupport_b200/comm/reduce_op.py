"""Reduction operator ids of the comm facade.  The numeric values follow torch.distributed / NCCL ordering so they can be
logged and compared across backends (reference: deepspeed/comm/reduce_op.py)."""
import enum

ReduceOp = enum.Enum("ReduceOp", ["SUM", "PRODUCT", "MIN", "MAX", "BAND", "BOR", "BXOR", "AVG", "UNUSED"], start=0)
ReduceOp.__doc__ = "SUM, PRODUCT, MIN, MAX, BAND, BOR, BXOR, AVG, UNUSED (values 0..8)"
