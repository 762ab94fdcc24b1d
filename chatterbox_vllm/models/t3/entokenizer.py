from chatterbox_vllm_b200.tokenizer import EnTokenizer  # noqa: F401
