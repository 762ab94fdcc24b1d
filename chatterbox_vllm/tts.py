from chatterbox_vllm_b200.tts import *  # noqa: F401,F403
from chatterbox_vllm_b200.tts import ChatterboxTTS, drop_invalid_tokens, guard_cap_tokens, postprocess_tokens, pre_cap_tokens  # noqa: F401
