"""Seeded inputs for the T3 glue fixtures — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

``oracle/make_golden.py::t3_glue`` feeds these to the REFERENCE'S OWN functions
(models/t3/t3.py: T3MultiModalProcessor.apply, T3VllmModel.split_prefill_decode /
get_input_embeddings / compute_logits, create_triangular_matrix — executed from
/root/reference with the vllm imports stubbed) and commits what they return as
tests/golden/t3_glue.npz; tests/test_oracle_pinning.py feeds the same inputs to
oracle/t3_ref.py, oracle/prompt_ref.py and chatterbox_vllm_b200/models/t3/processor.py
and requires the same bits.  Everything here is a pure function of its seed."""
from __future__ import annotations

import hashlib

import numpy as np
import torch

DIM = 1024
SPEECH_VOCAB = 8194
TEXT_VOCAB = {"en": 704, "mtl": 2454}
MARKERS = (695, 696, 697)


def tables(variant: str, seed: int = 1234):
    """fp32 embedding tables with the checkpoint's shapes (t3.py:323-331)."""
    g = torch.Generator().manual_seed(seed + (0 if variant == "en" else 1))
    n = lambda *s: torch.randn(*s, generator=g) * 0.02  # noqa: E731
    return dict(text_emb=n(TEXT_VOCAB[variant], DIM), speech_emb=n(SPEECH_VOCAB, DIM),
                text_pos=n(2048 + 2, DIM), speech_pos=n(4096 + 4, DIM))


def cond_emb(seed: int) -> torch.Tensor:
    return torch.randn(34, DIM, generator=torch.Generator().manual_seed(seed)) * 0.02


def text_ids(variant: str, n: int, seed: int) -> list[int]:
    """EN: [START]=255 … [STOP]=0; MTL: a language-tag id first and ids that exceed 1000 (SURVEY §8 d cfg 4).
    Never a marker id (695/696/697)."""
    rng = np.random.default_rng(seed)
    hi = TEXT_VOCAB[variant]
    body = [int(t) for t in rng.integers(3, hi, size=4 * n) if int(t) not in MARKERS and int(t) != 255]
    if variant == "mtl":
        body = body[:n - 1]
        body[0:2] = [708, 255]              # [en] tag then [START] (mtl_tts.py:380-401)
        body[2] = 2453                      # an id above SPEECH_TOKEN_OFFSET
        return body + [0]
    return [255] + body[:n - 2] + [0]


def prefix_cases():
    """(name, variant, n_text, ids seed, cond seed)."""
    return [("en_short", "en", 7, 11, 21), ("en_mid", "en", 23, 12, 22), ("mtl", "mtl", 12, 13, 23)]


def chunk_splits(P: int):
    """Cut points of the P-row prefix covering every slice case of SURVEY App. C: inside the cond rows, right
    after 695, at 696, right after 696, inside the text, before 697; two- and three-way."""
    two = [1, 5, 33, 34, 35, 34 + (P - 35) // 2, P - 2, P - 1]
    three = [(1, 34), (20, 40), (34, P - 1), (36, P - 3), (33, 35)]
    out = [(c,) for c in two if 0 < c < P]
    out += [t for t in three if 0 < t[0] < t[1] < P]
    return out


def digest(a) -> str:
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.tobytes()).hexdigest()


def sample(a, k: int = 4099):
    """A strided view small enough to commit, for readable failures next to the digest."""
    return np.ascontiguousarray(np.asarray(a).reshape(-1)[::k])


def hidden_rows(B: int, seed: int) -> torch.Tensor:
    """[B, 2048] fp32 'hidden states' (cond ‖ uncond) for compute_logits."""
    return torch.randn(B, 2 * DIM, generator=torch.Generator().manual_seed(seed))


def speech_head(seed: int = 77) -> torch.Tensor:
    return torch.randn(SPEECH_VOCAB, DIM, generator=torch.Generator().manual_seed(seed)) * 0.02 * 16


CFG_SCALES = (0.0, 0.5, 1.3)
