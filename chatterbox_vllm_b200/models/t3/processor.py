"""Prompt layout and input-embedding assembly of the T3 model class — the host-side contract between a
vLLM-style scheduler and the model (reference: models/t3/t3.py).

* ``T3MultiModalProcessor.apply`` (t3.py:160-296): ``final_prompt_ids = [695, ids[0] x 32, 696, *ids,
  697]`` plus the per-row side tensor ``[P, 1024]`` = cond_emb ‖ triangular position code ‖ zeros,
  the whole prompt marked as ONE multimodal placeholder.
* ``split_prefill_decode`` (t3.py:387-495): a flat ``input_ids`` batch is cut into prefill / decode
  segments by the markers 695 / 697 alone (multilingual text ids can exceed 1000).
* ``assemble_input_embeddings`` (t3.py:498-722): the eight prefix-slice cases.

Pure torch index arithmetic, device-agnostic (runs on CPU tensors in the CPU test-suite); the
engine's own prefill does the full-block case in one CUDA kernel (``cbx_prefill_embed``)."""
from __future__ import annotations

import os
import random
from typing import Mapping, Optional, Sequence, Union

import torch

PREFILL_COND_START_TOKEN = 695   # t3.py:38
PREFILL_COND_END_TOKEN = 696     # t3.py:39
PREFILL_END_TOKEN = 697          # t3.py:40
CONDITIONING_SIZE = 34           # t3.py:42
SPEECH_TOKEN_OFFSET = 1000       # t3.py:49
START_SPEECH_TOKEN = 6561        # t3_config.py


def _env_on(name: str) -> bool:
    return os.environ.get(name, "0").lower() in ("1", "true", "yes", "on")


def create_triangular_matrix(m: int, n: int) -> torch.Tensor:
    """[m, n] fp32 with row i = i+1 ones then zeros: encodes the text position of row i (t3.py:94-102)."""
    return (torch.arange(n).unsqueeze(0) <= torch.arange(m).unsqueeze(1)).float()


def build_prefill_inputs(prompt_ids: Sequence[int], conditionals: torch.Tensor) -> tuple[list[int], torch.Tensor]:
    """(final_prompt_ids, side tensor [P, width]) for one request (t3.py:229-262)."""
    prompt_ids = [int(t) for t in prompt_ids]
    assert conditionals.shape[0] == CONDITIONING_SIZE, "Conditionals must be CONDITIONING_SIZE tokens long"
    final_prompt_ids = ([PREFILL_COND_START_TOKEN] + [prompt_ids[0]] * (CONDITIONING_SIZE - 2) + [PREFILL_COND_END_TOKEN]
                        + prompt_ids + [PREFILL_END_TOKEN])
    width = conditionals.shape[1]
    side = torch.cat([conditionals,
                      create_triangular_matrix(len(prompt_ids), width).to(conditionals.device, conditionals.dtype),
                      torch.zeros(1, width, device=conditionals.device, dtype=conditionals.dtype)], dim=0)
    assert len(side) == len(final_prompt_ids), "Number of new conditionals does not match number of prompt ids"
    return final_prompt_ids, side


class T3MultiModalProcessor:
    """Stand-alone form of the reference's vLLM multimodal processor: same ``apply`` signature, returns
    the ``MultiModalInputs`` fields as a plain dict (``MultiModalInputs`` is a TypedDict in vLLM)."""

    def __init__(self, tokenizer, language_id: Optional[str] = None):
        self.tokenizer = tokenizer
        self.language_id = language_id

    def _tokenize(self, prompt: Union[str, list], tokenization_kwargs: Optional[Mapping[str, object]]) -> list[int]:
        if not isinstance(prompt, str):
            return [int(t) for t in prompt]
        lang = (tokenization_kwargs or {}).get("language_id", None) or self.language_id
        if isinstance(lang, str) and lang:                  # t3.py:138-143: the language tag travels inside the prompt
            tag = f"[{lang.lower()}]"
            if not prompt.startswith(tag):
                prompt = tag + prompt
        ids = self.tokenizer.encode(prompt)
        return [int(t) for t in ids]

    def apply(self, prompt: Union[str, list], mm_data: Mapping[str, object],
              hf_processor_mm_kwargs: Optional[Mapping[str, object]] = None,
              tokenization_kwargs: Optional[Mapping[str, object]] = None, return_mm_hashes: bool = False) -> dict:
        prompt_ids = self._tokenize(prompt, tokenization_kwargs)
        if _env_on("CHATTERBOX_STRIP_LLM_EOS"):            # t3.py:209-221 (off by default: id 2 is a valid text token)
            prompt_ids = [t for t in prompt_ids if t != 2]
        conditionals = mm_data.get("conditionals", None)
        assert conditionals is not None and len(conditionals) > 0, "Conditionals are required for prefill"
        if isinstance(conditionals, torch.Tensor) and conditionals.dim() == 2:
            conditionals = [conditionals]
        assert len(conditionals) == 1, "Only one conditional embedding is supported for prefill"
        final_prompt_ids, side = build_prefill_inputs(prompt_ids, conditionals[0])
        return {
            "type": "multimodal",
            "prompt": prompt,
            "prompt_token_ids": final_prompt_ids,
            "mm_kwargs": {"conditionals": [side]},
            "mm_hashes": {"conditionals": [str(random.random())]},     # the reference does not hash (t3.py:286-289)
            # the conditionals "modify" the entire prompt, so the side tensor reaches every prefix row (t3.py:290-295)
            "mm_placeholders": {"conditionals": [{"offset": 0, "length": len(final_prompt_ids), "is_embed": None}]},
        }


def split_prefill_decode(input_ids: torch.Tensor, multimodal_embeddings: Sequence[torch.Tensor]):
    """[(ids, side rows | None)]: prefill segments carry their rows of the concatenated side tensors,
    decode segments carry None (t3.py:387-495)."""
    if len(input_ids) == 0:
        return []
    remaining = torch.cat(list(multimodal_embeddings), dim=0)
    strip = _env_on("CHATTERBOX_STRIP_LLM_EOS")
    out, buf = [], []
    in_prefill: Optional[bool] = None                       # undecided: assume prefill until the first 697

    def flush(is_prefill: bool):
        nonlocal buf, remaining
        if not buf:
            return
        seg = torch.tensor(buf, device=input_ids.device)
        if is_prefill:
            take = min(len(buf), len(remaining))
            mme, remaining = remaining[:take], remaining[take:]
            out.append((seg, mme))
        else:
            out.append((seg, None))
        buf = []

    for tok in input_ids.tolist():
        if strip and tok == 2:
            continue
        if tok == PREFILL_COND_START_TOKEN:
            if in_prefill is not None:
                flush(in_prefill)
            in_prefill = True
            buf.append(tok)
        elif tok == PREFILL_END_TOKEN:
            buf.append(tok)
            flush(True)
            in_prefill = False
        else:
            if in_prefill is None:
                in_prefill = True
            buf.append(tok)
    if buf:
        flush(bool(in_prefill))
    return out


class EmbeddingTables:
    """The four tables the assembly reads, in the model dtype (bf16 on device)."""

    def __init__(self, text_emb, text_pos, speech_emb, speech_pos):
        self.text_emb, self.text_pos, self.speech_emb, self.speech_pos = text_emb, text_pos, speech_emb, speech_pos
        self.dim = text_emb.shape[1]
        self.dtype = text_emb.dtype


def _positions_from_side(side: torch.Tensor, n: int, limit: int):
    """Text positions of a mid-prefix slice = ones per row of the triangular code, minus 1
    (t3.py:631-633,668-669).  The reference indexes its position table with ``text_pos.tolist()`` — a
    list of Python FLOATS, which torch truncates toward zero and wraps when negative — so a slice whose
    side rows are not a position code (cond rows reached without their 695 marker) still yields rows:
    ``text_emb[id] + text_pos[trunc(rowsum - 1)]``.  Reproduced as executed (pinned by
    tests/golden/t3_glue.npz).  None when an index falls outside the table (the reference then raises
    inside its try block and falls back to zeros)."""
    if len(side) < n:
        return None
    pos = (side[:n].float().sum(dim=1) - 1).trunc().long()
    pos = torch.where(pos < 0, pos + limit, pos)
    if n and (int(pos.min()) < 0 or int(pos.max()) >= limit):
        return None
    return pos


def assemble_input_embeddings(input_ids: torch.Tensor, multimodal_embeddings, tb: EmbeddingTables) -> torch.Tensor:
    """[N, 2·dim] = cond half ‖ uncond half for a flat batch of prefill slices and decode tokens
    (t3.py:498-722; the uncond half sees zeros instead of text embeddings, t3.py:563)."""
    dev, dt, dim = input_ids.device, tb.dtype, tb.dim
    if multimodal_embeddings is None or len(multimodal_embeddings) == 0:
        e = tb.speech_emb[input_ids - SPEECH_TOKEN_OFFSET]            # pure decode (t3.py:503-514)
        return torch.cat([e, e], dim=1)
    sos = (tb.speech_emb[START_SPEECH_TOKEN] + tb.speech_pos[0]).unsqueeze(0)

    def text_rows(ids, pos=None):
        if pos is None:
            pos = torch.arange(len(ids), device=dev)
        return tb.text_emb[ids] + tb.text_pos[pos]

    def pair(cond_parts, uncond_parts):
        return torch.cat([torch.cat(cond_parts, dim=0), torch.cat(uncond_parts, dim=0)], dim=1)

    out = []
    for ids, side in split_prefill_decode(input_ids, multimodal_embeddings):
        n = len(ids)
        if side is None:                                                # decode tokens inside a mixed batch
            e = tb.speech_emb[ids - SPEECH_TOKEN_OFFSET]
            out.append(torch.cat([e, e], dim=1))
            continue
        side = side.to(device=dev)
        first, last = int(ids[0]), int(ids[-1])
        zeros = torch.zeros(n, 2 * dim, device=dev, dtype=dt)
        if first == PREFILL_COND_START_TOKEN and last == PREFILL_END_TOKEN:        # full block (t3.py:548-569)
            text = text_rows(ids[CONDITIONING_SIZE:-1])
            cond = side[:CONDITIONING_SIZE].to(dt)
            out.append(pair([cond, text, sos], [cond, torch.zeros_like(text), sos]))
        elif first == PREFILL_COND_START_TOKEN and n > 1:                          # head of a split block (t3.py:570-592)
            text = text_rows(ids[CONDITIONING_SIZE:])
            cond = side[:min(CONDITIONING_SIZE, len(side))].to(dt)
            e = pair([cond, text], [cond, torch.zeros_like(text)])
            out.append(e if len(e) == n else zeros)
        elif last == PREFILL_END_TOKEN:                                            # tail of a split block
            idx = torch.where(ids == PREFILL_COND_END_TOKEN)[0]
            if len(idx) > 0:                                                       # ... with the end of the cond rows (t3.py:604-623)
                k = int(idx[0])
                text = text_rows(ids[k + 1:-1])
                cond = side[:k + 1].to(dt)
                out.append(pair([cond, text, sos], [cond, torch.zeros_like(text), sos]))
            else:                                                                  # ... text tail only (t3.py:624-644)
                pos = _positions_from_side(side, n - 1, len(tb.text_pos))
                if pos is None:
                    out.append(zeros)
                    continue
                try:                                                               # (the reference has no guard on this branch
                    text = text_rows(ids[:-1], pos)                                # and would raise into the engine: zeros here)
                except IndexError:
                    out.append(zeros)
                    continue
                out.append(pair([text, sos], [torch.zeros_like(text), sos]))
        elif n == 1 and first in (PREFILL_COND_END_TOKEN, PREFILL_COND_START_TOKEN):   # lone 696 / 695 (t3.py:646-660)
            m = side.to(dt)
            out.append(torch.cat([m, m], dim=1))
        elif first != PREFILL_COND_START_TOKEN and last != PREFILL_END_TOKEN:      # middle text slice (t3.py:661-685)
            pos = _positions_from_side(side, n, len(tb.text_pos))
            if pos is None:
                out.append(zeros)
                continue
            try:                                                                   # the reference's try/except (t3.py:666-685):
                text = text_rows(ids, pos)                                         # an id outside the text table -> zeros
            except IndexError:
                out.append(zeros)
                continue
            out.append(torch.cat([text, torch.zeros_like(text)], dim=1))
        else:                                                                      # unknown: keep alignment (t3.py:686-700)
            out.append(zeros)
    if len(out) == 0:                                                              # t3.py:702-714
        return torch.zeros(len(input_ids), 2 * dim, device=dev, dtype=dt)
    return torch.cat(out, dim=0)
