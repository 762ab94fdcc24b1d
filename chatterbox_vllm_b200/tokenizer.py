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


# ---------------------------------------------------------------------------------------------------
# Multilingual text tokenizer (reference: src/chatterbox_vllm/models/t3/mtl_tokenizer.py:223-476).
# Same pipeline, without the HF PreTrainedTokenizer shim: leading "[xx]" language tag detected and
# stripped, lower-casing + NFKD (with [START]/[STOP] restored to canonical casing), language-specific
# normalisation, the "[xx]" tag put back in front, spaces -> [SPACE], BPE of the checkpoint's
# grapheme_mtl_merged_expanded_v1.json / mtl_tokenizer.json (2454 ids).  The language normalisers that
# need third-party packages (pykakasi for ja, dicta_onnx for he, russian_text_stresser for ru,
# spacy_pkuseg for zh segmentation) are optional exactly as in the reference: absent -> the text passes
# through unchanged with a warning.  Korean Jamo decomposition and the Cangjie mapping (from the
# checkpoint's Cangjie5_TC.json when it lies next to the vocabulary) are implemented here.
import logging as _logging
import re as _re
from unicodedata import category as _category, normalize as _normalize

from .text_utils import SUPPORTED_LANGUAGES

_log = _logging.getLogger(__name__)


def korean_normalize(text: str) -> str:
    """Hangul syllables -> initial / medial / final Jamo (mtl_tokenizer.py:101-122)."""
    out = []
    for ch in text:
        if "가" <= ch <= "힯":
            base = ord(ch) - 0xAC00
            out.append(chr(0x1100 + base // (21 * 28)) + chr(0x1161 + (base % (21 * 28)) // 28)
                       + (chr(0x11A7 + base % 28) if base % 28 > 0 else ""))
        else:
            out.append(ch)
    return "".join(out).strip()


def _optional(normaliser_name: str, fn):
    def run(text: str) -> str:
        try:
            return fn(text)
        except ImportError:
            _log.warning("%s not available - text passed through unchanged", normaliser_name)
            return text
        except Exception as e:                            # the reference swallows normaliser failures too
            _log.warning("%s failed: %s", normaliser_name, e)
            return text
    return run


def _hiragana(text: str) -> str:                          # mtl_tokenizer.py:44-83
    import pykakasi
    out = []
    for r in pykakasi.kakasi().convert(text):
        inp, hira = r["orig"], r["hira"]
        if any(19968 <= ord(c) <= 40959 for c in inp):
            out.append((" " + hira) if hira and hira[0] in ("は", "へ") else hira)
        else:
            out.append(inp)
    return _normalize("NFKD", "".join(out))


def _hebrew(text: str) -> str:                            # mtl_tokenizer.py:86-98
    from dicta_onnx import Dicta
    return Dicta().add_diacritics(text)


def _russian(text: str) -> str:                           # mtl_tokenizer.py:204-220
    from russian_text_stresser.text_stresser import RussianTextStresser
    return RussianTextStresser().stress_text(text)


class CangjieConverter:
    """Chinese glyphs -> "[cj_x]…[cj_.]" Cangjie code tokens (mtl_tokenizer.py:125-201).  The mapping comes from
    Cangjie5_TC.json (lines "glyph<TAB>code"); without it glyphs pass through unchanged."""

    def __init__(self, mapping_file: str | None):
        self.word2cj, self.cj2word, self.segmenter = {}, {}, None
        if mapping_file and os.path.isfile(mapping_file):
            import json
            for entry in json.load(open(mapping_file, encoding="utf-8")):
                word, code = entry.split("\t")[:2]
                self.word2cj[word] = code
                self.cj2word.setdefault(code, []).append(word)
        else:
            _log.warning("Cangjie5_TC.json not found - Chinese glyphs are not converted")
        try:
            from spacy_pkuseg import pkuseg
            self.segmenter = pkuseg()
        except ImportError:
            _log.warning("pkuseg not available - Chinese segmentation will be skipped")

    def __call__(self, text: str) -> str:
        if self.segmenter is not None:
            text = " ".join(self.segmenter.cut(text))
        out = []
        for ch in text:
            code = self.word2cj.get(ch) if _category(ch) == "Lo" else None
            if code is None:
                out.append(ch)
                continue
            idx = self.cj2word[code].index(ch)
            out.append("".join(f"[cj_{c}]" for c in code + (str(idx) if idx > 0 else "")) + "[cj_.]")
        return "".join(out)


class MTLTokenizer:
    def __init__(self, vocab_file: str):
        from tokenizers import Tokenizer
        self.tokenizer = Tokenizer.from_file(vocab_file)
        voc = self.tokenizer.get_vocab()
        if SOT not in voc or EOT not in voc:
            raise ValueError(f"{vocab_file}: vocabulary lacks {SOT}/{EOT}")
        self.cangjie = CangjieConverter(os.path.join(os.path.dirname(os.path.abspath(vocab_file)), "Cangjie5_TC.json"))
        self.normalisers = {"zh": self.cangjie, "ja": _optional("pykakasi", _hiragana), "he": _optional("dicta_onnx", _hebrew),
                            "ko": korean_normalize, "ru": _optional("russian_text_stresser", _russian)}

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path: str | None = None, **_):
        """Resolution order of the reference (mtl_tokenizer.py:255-293): the given directory, then
        ./t3-multilingual-model; in each the expanded grapheme vocabulary first, then mtl_tokenizer.json."""
        dirs = [d for d in (pretrained_model_name_or_path, os.path.join(os.getcwd(), "t3-multilingual-model"))
                if d and os.path.isdir(d)]
        cands = [os.path.join(d, f) for d in dirs for f in ("grapheme_mtl_merged_expanded_v1.json", "mtl_tokenizer.json")]
        for c in cands:
            if os.path.isfile(c):
                return cls(c)
        raise FileNotFoundError("MTLTokenizer: could not locate a tokenizer JSON (grapheme_mtl_merged_expanded_v1.json or "
                                "mtl_tokenizer.json); looked in: " + ", ".join(cands))

    def get_vocab(self) -> dict:
        return self.tokenizer.get_vocab()

    @property
    def vocab_size(self) -> int:
        return len(self.tokenizer.get_vocab())

    def preprocess_text(self, text: str, lowercase: bool = True, nfkd_normalize: bool = True) -> str:
        if lowercase:
            text = text.lower()
        if nfkd_normalize:
            text = _normalize("NFKD", text)
        return _re.sub(r"\[stop\]", EOT, _re.sub(r"\[start\]", SOT, text))

    def prepare(self, text: str, language_id: str | None = None, lowercase: bool = True, nfkd_normalize: bool = True) -> str:
        """The string that reaches the BPE (mtl_tokenizer.py:455-466)."""
        if not language_id and text.startswith("[") and "]" in text[:8]:
            tag = text[1:text.index("]")].lower()
            if tag in SUPPORTED_LANGUAGES:                     # only a supported code counts as a language tag
                language_id, text = tag, text[text.index("]") + 1:]
        elif not language_id and text.startswith("<") and ">" in text[:8]:
            # tts.py:330-333 (the English class on the multilingual checkpoint) prefixes "<xx>" and says "this will be
            # parsed and replaced in the tokenizer"; the reference's tokenizer never does.  Parsed here, like "[xx]".
            tag = text[1:text.index(">")].lower()
            if tag in SUPPORTED_LANGUAGES:
                language_id, text = tag, text[text.index(">") + 1:]
        text = self.preprocess_text(text, lowercase, nfkd_normalize)
        if language_id in self.normalisers:
            text = self.normalisers[language_id](text)
        if language_id:
            tag = f"[{language_id.lower()}]"
            if not text.startswith(tag):
                text = tag + text
        return text.replace(" ", SPACE)

    def encode(self, text: str, language_id: str | None = None, lowercase: bool = True, nfkd_normalize: bool = True) -> list[int]:
        return self.tokenizer.encode(self.prepare(text, language_id, lowercase, nfkd_normalize)).ids

    def tokenize(self, text: str, language_id: str | None = None) -> list[str]:
        return self.tokenizer.encode(self.prepare(text, language_id)).tokens

    def decode(self, ids) -> str:
        txt = self.tokenizer.decode(list(ids), skip_special_tokens=False)
        return txt.replace(" ", "").replace(SPACE, " ").replace(EOT, "").replace(UNK, "")
