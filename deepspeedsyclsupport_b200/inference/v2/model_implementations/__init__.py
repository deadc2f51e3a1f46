from .ragged_transformer import ArchSpec, RaggedTransformer, SPEC_FROM_HF  # noqa: F401
from .inference_model_base import DSInferenceModelBase  # noqa: F401
from .inference_transformer_base import DSMoETransformerModelBase, DSTransformerModelBase  # noqa: F401
from .inference_policy_base import POLICIES, ContainerMap, InferenceV2Policy, policy_for  # noqa: F401
from .layer_container_base import LayerContainer  # noqa: F401
from .parameter_base import ParameterBase, ParamList  # noqa: F401
from . import falcon, gpt2, gpt_neox, gptj, llama_v2, mistral, mixtral, opt, phi  # noqa: F401  (registers the policies)
from .falcon import FalconPolicy  # noqa: F401
from .llama_v2 import Llama2Policy  # noqa: F401
from .mistral import MistralPolicy  # noqa: F401
from .mixtral import MixtralPolicy  # noqa: F401
from .opt import OPTPolicy  # noqa: F401
from .phi import PhiPolicy  # noqa: F401

FAMILIES = ("llama_v2", "mistral", "mixtral", "opt", "falcon", "phi", "gpt2", "gptj", "gpt_neox")
