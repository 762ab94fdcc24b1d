from chatterbox_vllm_b200.text_utils import SUPPORTED_LANGUAGES, punc_norm  # noqa: F401
