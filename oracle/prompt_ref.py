"""CPU oracle for the host-side prompt layout / pre-cap / token post-processing —
TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates (reference file:line):
  models/t3/t3.py:38-49           marker ids, CONDITIONING_SIZE, SPEECH_TOKEN_OFFSET
  models/t3/t3.py:94-102,229-262  final_prompt_ids + the per-row side tensor
  text_utils.py:2-41              punc_norm
  tts.py:327                      "[START]" + punc_norm(p) + "[STOP]"
  tts.py:350-362                  sampling clamps / CHATTERBOX_DETERMINISTIC
  tts.py:372-399                  max_tokens pre-cap (min over the batch)
  tts.py:448-465                  stop cut, tail crop, de-offset, < 6561 filter
  models/s3tokenizer/__init__.py:38-78  drop_invalid_tokens

PARITY STATUS: the reference has no tests; final_prompt_ids / side_tensor are pinned bit for bit
to the reference's own T3MultiModalProcessor.apply executed from /root/reference
(tests/golden/t3_glue.npz, tests/test_reference_glue.py); punc_norm and the chunker to the
reference's own text_utils.py / benchmark.py (tests/golden/prompt_lengths.json), and the layout
arithmetic to the README's log line (28 754 prompt tokens for the 154 chunks of
docs/benchmark-text-1.txt — README.md:151).  The pre-cap / post-processing functions restate
tts.py and are unpinned (tts.py needs librosa / s3tokenizer / diffusers to import).
"""
from __future__ import annotations

import math
import re

import numpy as np

PREFILL_COND_START_TOKEN = 695
PREFILL_COND_END_TOKEN = 696
PREFILL_END_TOKEN = 697
CONDITIONING_SIZE = 34
SPEECH_TOKEN_OFFSET = 1000
START_SPEECH, STOP_SPEECH = 6561, 6562


def punc_norm(text: str) -> str:
    if len(text) == 0:
        return "You need to add some text for me to talk."
    if text[0].islower():
        text = text[0].upper() + text[1:]
    text = " ".join(text.split())
    for old, new in [("...", ", "), ("…", ", "), (":", ","), (" - ", ", "), (";", ", "),
                     ("—", "-"), ("–", "-"), (" ,", ","), ("“", "\""), ("”", "\""),
                     ("‘", "'"), ("’", "'")]:
        text = text.replace(old, new)
    text = text.rstrip(" ")
    if not any(text.endswith(p) for p in {".", "!", "?", "-", ",", "、", "，", "。", "？", "！"}):
        text += "."
    return text


def wrap_prompt(p: str, language_id: str | None = None, multilingual: bool = False) -> str:
    s = "[START]" + punc_norm(p) + "[STOP]"
    if multilingual:
        s = f"<{language_id.lower()}>{s}"
    return s


def final_prompt_ids(text_ids: list[int]) -> list[int]:
    """t3.py:229-240: [695, ids[0]×32, 696, *ids, 697]  (P = 34 + n_text + 1)."""
    return ([PREFILL_COND_START_TOKEN] + [text_ids[0]] * (CONDITIONING_SIZE - 2)
            + [PREFILL_COND_END_TOKEN] + list(text_ids) + [PREFILL_END_TOKEN])


def side_tensor(cond_emb: np.ndarray, n_text: int) -> np.ndarray:
    """t3.py:252-262: [P, 1024] fp32 = cond_emb ‖ tri(n_text, 1024) ‖ zeros[1]."""
    width = cond_emb.shape[1]
    tri = (np.arange(width)[None, :] <= np.arange(n_text)[:, None]).astype(np.float32)
    return np.concatenate([cond_emb.astype(np.float32), tri, np.zeros((1, width), np.float32)], 0)


def clamp_sampling(temperature: float, top_p: float, language_id, deterministic: bool):
    """tts.py:350-362."""
    if language_id:
        temperature, top_p = min(temperature, 0.5), min(top_p, 0.5)
    if deterministic:
        temperature, top_p = 0.0, 1.0
    return temperature, top_p


def pre_cap_tokens(prompts: list[str], max_tokens: int, max_model_len: int, *, tpc: float = 2.2,
                   tmin: int = 64, tmax: int = 1200, guard_mult: float = 1.6, margin: int = 16) -> int:
    """tts.py:372-399 (prompts are the wrapped strings; bracket tags are stripped)."""
    caps = []
    for p in prompts:
        clean = re.sub(r"\[[^\]]+\]", "", p)
        clen = sum(1 for c in clean if not c.isspace())
        est = max(tmin, min(tmax, int(math.ceil(clen * tpc))))
        caps.append(int(math.ceil(est * guard_mult)))
    cap = max(1, min(caps) + max(0, margin)) if caps else min(max_tokens, max_model_len)
    return min(cap, min(max_tokens, max_model_len))


def drop_invalid_tokens(x: np.ndarray) -> np.ndarray:
    """s3tokenizer/__init__.py:38-78: slice (first SOS+1 .. first EOS) when present."""
    x = np.asarray(x)
    sos = np.nonzero(x == START_SPEECH)[0]
    eos = np.nonzero(x == STOP_SPEECH)[0]
    s = int(sos[0]) + 1 if len(sos) else 0
    e = int(eos[0]) if len(eos) else len(x)
    return x[s:e]


def postprocess_tokens(tok_ids: list[int], tail_crop_k: int = 2) -> np.ndarray:
    """tts.py:448-465 on vLLM-space ids (speech id + 1000)."""
    stop = STOP_SPEECH + SPEECH_TOKEN_OFFSET
    if stop in tok_ids:
        tok_ids = tok_ids[:max(0, tok_ids.index(stop) - max(0, int(tail_crop_k)))]
    x = np.asarray([t - SPEECH_TOKEN_OFFSET for t in tok_ids], np.int64)
    x = drop_invalid_tokens(x)
    return x[x < 6561]
