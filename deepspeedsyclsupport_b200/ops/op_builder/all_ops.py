"""Builder registry for `XBuilder().load()` callers (parity: op_builder/all_ops.py:17-32 and each op_builder/*.py).
Every builder resolves to the same two in-tree libraries; `load()` returns a light module-like object exposing the
library handle so user code written against the reference (`FusedAdamBuilder().load()`) keeps working."""
from .. import native


class _LoadedOp:

    def __init__(self, name, lib):
        self.name, self.lib = name, lib

    def __getattr__(self, item):
        return getattr(self.lib, item)


class OpBuilder:
    NAME = "op"
    HOST = False

    def is_compatible(self, verbose=True):
        return True

    def absolute_name(self):
        return f"deepspeedsyclsupport_b200.ops.{self.NAME}_op"

    def sources(self):
        from . import builder
        return builder.host_sources() if self.HOST else builder.device_sources()

    def load(self, verbose=True):
        return _LoadedOp(self.NAME, native.host_lib() if self.HOST else native.lib())

    def builder(self):
        return self


def _mk(cls_name, name, host=False):
    return type(cls_name, (OpBuilder,), {"NAME": name, "HOST": host})


BUILDERS = {n: _mk(n, nm, h) for n, nm, h in [
    ("FusedAdamBuilder", "fused_adam", False), ("FusedLambBuilder", "fused_lamb", False),
    ("FusedLionBuilder", "fused_lion", False), ("CPUAdamBuilder", "cpu_adam", True),
    ("CPUAdagradBuilder", "cpu_adagrad", True), ("CPULionBuilder", "cpu_lion", True),
    ("QuantizerBuilder", "quantizer", False), ("AsyncIOBuilder", "async_io", True),
    ("TransformerBuilder", "transformer", False), ("StochasticTransformerBuilder", "stochastic_transformer", False),
    ("InferenceBuilder", "transformer_inference", False), ("InferenceCoreBuilder", "inference_core_ops", False),
    ("RaggedOpsBuilder", "ragged_device_ops", False), ("RaggedUtilsBuilder", "ragged_ops", True),
    ("InferenceCutlassBuilder", "cutlass_ops", False), ("RandomLTDBuilder", "random_ltd", False),
    ("SparseAttnBuilder", "sparse_attn", True), ("SpatialInferenceBuilder", "spatial_inference", False),
    ("EvoformerAttnBuilder", "evoformer_attn", False), ("CCLCommBuilder", "deepspeed_ccl_comm", True),
    ("UtilsBuilder", "utils", True),
]}
ALL_OPS = {b.NAME: b() for b in BUILDERS.values()}
globals().update(BUILDERS)
