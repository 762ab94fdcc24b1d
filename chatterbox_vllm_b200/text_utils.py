"""Host-side text normalisation on the request path (reference: src/chatterbox_vllm/text_utils.py:2-41
``punc_norm`` and :44-68 ``SUPPORTED_LANGUAGES``).  Behaviour-identical restatement; checked against
the oracle's copy of the rules in tests/test_host_surface.py."""
from __future__ import annotations

_EMPTY_PROMPT = "You need to add some text for me to talk."
# (needle, replacement) applied in this order — order matters ("..." before ":" etc.)
_PUNC_RULES = (
    ("...", ", "), ("…", ", "), (":", ","), (" - ", ", "), (";", ", "), ("—", "-"), ("–", "-"),
    (" ,", ","), ("“", '"'), ("”", '"'), ("‘", "'"), ("’", "'"),
)
_ENDERS = (".", "!", "?", "-", ",", "、", "，", "。", "？", "！")

SUPPORTED_LANGUAGES = {
    "ar": "Arabic", "da": "Danish", "de": "German", "el": "Greek", "en": "English", "es": "Spanish",
    "fi": "Finnish", "fr": "French", "he": "Hebrew", "hi": "Hindi", "it": "Italian", "ja": "Japanese",
    "ko": "Korean", "ms": "Malay", "nl": "Dutch", "no": "Norwegian", "pl": "Polish", "pt": "Portuguese",
    "ru": "Russian", "sv": "Swedish", "sw": "Swahili", "tr": "Turkish", "zh": "Chinese",
}


def punc_norm(text: str) -> str:
    if not text:
        return _EMPTY_PROMPT
    head = text[0]
    if head.islower():
        text = head.upper() + text[1:]
    text = " ".join(text.split())                 # collapse every whitespace run to one space
    for needle, repl in _PUNC_RULES:
        if needle in text:
            text = text.replace(needle, repl)
    text = text.rstrip(" ")
    return text if text.endswith(_ENDERS) else text + "."
