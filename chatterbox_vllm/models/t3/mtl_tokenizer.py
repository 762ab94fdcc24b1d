from chatterbox_vllm_b200.tokenizer import MTLTokenizer  # noqa: F401
