"""``chatterbox_vllm.models.t3`` (reference: models/t3/__init__.py:1-7): the vLLM-facing model class, its processor,
SPEECH_TOKEN_OFFSET, and the registration the reference performs at import — done here when vLLM is importable."""
from chatterbox_vllm_b200.models.t3 import *  # noqa: F401,F403
from chatterbox_vllm_b200.models.t3 import SPEECH_TOKEN_OFFSET, register_with_vllm  # noqa: F401


def __getattr__(name):
    if name == "T3VllmModel":
        from chatterbox_vllm_b200.models.t3 import T3VllmModel
        return T3VllmModel
    raise AttributeError(name)
