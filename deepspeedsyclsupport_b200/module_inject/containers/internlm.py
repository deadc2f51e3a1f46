"""InternLM decoder layer (parity: deepspeed/module_inject/containers/internlm.py): the Llama layout with biases on the
attention projections; the class is found by name because the model ships as remote code."""
import importlib

from .llama import DS_LLAMAContainer, LLAMALayerPolicy


class DS_InternLMContainer(DS_LLAMAContainer):
    pass


class InternLMLayerPolicy(LLAMALayerPolicy):
    _orig_layer_class = []
    _orig_layer_class_inited = False

    def __init__(self, client_module, inference=True):
        super().__init__(client_module, inference)
        self._init_orig_layer_class_once()

    def _init_orig_layer_class_once(self):
        if InternLMLayerPolicy._orig_layer_class_inited:
            return
        for sub_pkg in ("", ".internlm-7b", ".internlm-chat-7b"):
            try:
                mod = importlib.import_module(f"transformers_modules{sub_pkg}.modeling_internlm")
                if mod.InternLMDecoderLayer not in InternLMLayerPolicy._orig_layer_class:
                    InternLMLayerPolicy._orig_layer_class.append(mod.InternLMDecoderLayer)
            except ImportError:
                continue
        InternLMLayerPolicy._orig_layer_class_inited = True
