"""English text tokenizer of the T3 path (reference: src/chatterbox_vllm/models/t3/entokenizer.py:19-117).

Same rules — spaces become the ``[SPACE]`` token, then the ``tokenizers`` BPE of the checkpoint's
``tokenizer.json`` (704 ids; ``[START]``=255, ``[STOP]``=0) — without the HF ``PreTrainedTokenizer``
shim vLLM needed.  The vocabulary file is resolved like the reference does (entokenizer.py:47-83):
an explicit directory, then ``./t3-model/tokenizer.json``."""
from __future__ import annotations

import os

SOT, EOT, UNK, SPACE = "[START]", "[STOP]", "[UNK]", "[SPACE]"


class EnTokenizer:
    def __init__(self, vocab_file: str):
        from tokenizers import Tokenizer
        self.tokenizer = Tokenizer.from_file(vocab_file)
        voc = self.tokenizer.get_vocab()
        if SOT not in voc or EOT not in voc:
            raise ValueError(f"{vocab_file}: vocabulary lacks {SOT}/{EOT}")

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path: str | None = None, **_):
        cands = []
        if pretrained_model_name_or_path and os.path.isdir(pretrained_model_name_or_path):
            cands.append(os.path.join(pretrained_model_name_or_path, "tokenizer.json"))
        cands.append(os.path.join(os.getcwd(), "t3-model", "tokenizer.json"))
        for c in cands:
            if os.path.isfile(c):
                return cls(c)
        raise FileNotFoundError("EnTokenizer: could not locate tokenizer.json (looked in: %s)" % ", ".join(cands))

    def get_vocab(self) -> dict:
        return self.tokenizer.get_vocab()

    @property
    def vocab_size(self) -> int:
        return self.tokenizer.get_vocab_size()

    def tokenize(self, text: str) -> list[str]:
        return self.tokenizer.encode(text.replace(" ", SPACE)).tokens

    def encode(self, text: str) -> list[int]:
        """Ids of ``text`` (no special tokens added — the caller wraps with [START]/[STOP], tts.py:327)."""
        return self.tokenizer.encode(text.replace(" ", SPACE)).ids

    def decode(self, ids: list[int]) -> str:
        toks = [self.tokenizer.id_to_token(i) or UNK for i in ids]
        s = "".join(toks).replace(" ", "").replace(SPACE, " ")
        return s.replace(EOT, "").replace(UNK, "")
