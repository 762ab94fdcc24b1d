"""Drop-in alias: the reference's import paths (``from chatterbox_vllm.tts import ChatterboxTTS``, example-tts.py:1-8;
``chatterbox_vllm.mtl_tts``, ``chatterbox_vllm.models.t3``, ``chatterbox_vllm.text_utils``) resolve to the B200
implementation in ``chatterbox_vllm_b200``.  Nothing lives here but the names."""
from chatterbox_vllm_b200 import *  # noqa: F401,F403
