from chatterbox_vllm_b200.mtl_tts import ChatterboxMultilingualTTS  # noqa: F401
from chatterbox_vllm_b200.text_utils import SUPPORTED_LANGUAGES  # noqa: F401
