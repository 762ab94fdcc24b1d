"""The vLLM-facing T3 model class and its prompt processor (reference: models/t3/__init__.py:5-7,
models/t3/t3.py).  Importing this package never imports vLLM; ``register_with_vllm()`` does, and
registers the class under the reference's architecture name."""
from .processor import (CONDITIONING_SIZE, PREFILL_COND_END_TOKEN, PREFILL_COND_START_TOKEN, PREFILL_END_TOKEN,
                        SPEECH_TOKEN_OFFSET, EmbeddingTables, T3MultiModalProcessor, assemble_input_embeddings,
                        build_prefill_inputs, create_triangular_matrix, split_prefill_decode)

__all__ = ["T3VllmModel", "T3MultiModalProcessor", "register_with_vllm", "build_prefill_inputs", "split_prefill_decode",
           "assemble_input_embeddings", "EmbeddingTables", "create_triangular_matrix", "CONDITIONING_SIZE",
           "SPEECH_TOKEN_OFFSET", "PREFILL_COND_START_TOKEN", "PREFILL_COND_END_TOKEN", "PREFILL_END_TOKEN"]


def __getattr__(name):                                   # T3VllmModel needs the CUDA library: import it lazily
    if name == "T3VllmModel":
        from .t3 import T3VllmModel
        return T3VllmModel
    raise AttributeError(name)


def register_with_vllm() -> bool:
    """ModelRegistry.register_model("ChatterboxT3", T3VllmModel) + the EnTokenizer registration the
    reference performs at import (models/t3/__init__.py:5-7).  Returns False when vLLM is absent.
    The class runs on its own KV cache (see t3.py), so under a stock vLLM it is the numerical
    drop-in for embeddings / forward / logits, not a user of vLLM's attention backends."""
    try:
        from vllm import ModelRegistry
    except Exception:
        return False
    ModelRegistry.register_model("ChatterboxT3", "chatterbox_vllm_b200.models.t3.t3:T3VllmModel")
    try:                                                 # the registry moved between vLLM releases
        from vllm.transformers_utils.tokenizer_base import TokenizerRegistry
        TokenizerRegistry.register("EnTokenizer", "chatterbox_vllm_b200.tokenizer", "EnTokenizer")
    except Exception:
        pass
    return True
