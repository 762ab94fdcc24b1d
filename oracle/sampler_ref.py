"""CPU oracle for the fused CFG-sampler — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates, in numpy fp32, the vLLM v1 sampler chain the reference reaches through
``SamplingParams`` (tts.py:408-418), in the order vLLM applies it
(vllm-0.22 v1/sample/sampler.py:67-300, 366-403):

  repetition penalty   model_executor/layers/utils.py:51-89, _custom_ops.py:360-373
  greedy argmax        sampler.py:235-236  (on the PENALISED logits, if T < 1e-5)
  temperature          sampler.py:223-233
  min-p                v1/sample/logits_processor/builtin.py:101-112
  top-k / top-p        v1/sample/ops/topk_topp_sampler.py:328-369
  exponential race     topk_topp_sampler.py:395-416   argmax(softmax(l) / Exp(1))

Two things are *specified here* because the reference leaves them to library
accident: (1) ties — equal logits are ranked by ascending token id (lower id
wins argmax and is kept first by top-p; top-k keeps every tie with the k-th value, as vLLM); (2) the random stream — torch's
global Philox state is replaced by a counter-based Philox4x32-10 keyed by
(seed, request uid, step, token id), so results do not depend on batch layout or
on how many GPUs the requests are sharded over.

PARITY STATUS: unpinned by the reference (no tests there); pinned against vLLM
0.22's CPU-runnable torch implementations in tests/test_oracle_pinning.py.
"""
from __future__ import annotations

import numpy as np

SAMPLING_EPS = 1e-5          # sampler.py:17
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """Philox-4x32-10 (Salmon et al., SC'11).  ctr: 4 uint32 arrays (broadcastable),
    key: 2 python ints.  Returns 4 uint32 arrays."""
    c = [np.asarray(x, dtype=np.uint64) & _MASK for x in ctr]
    c = list(np.broadcast_arrays(*c))
    k0, k1 = key[0] & 0xFFFFFFFF, key[1] & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c[0]
        p1 = _M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return [x.astype(np.uint32) for x in c]


def exp1_variates(seed: int, uid: int, step: int, token_ids) -> np.ndarray:
    """Exp(1) draws e_i = −ln(u_i), u_i = ((x_i >> 9) + 0.5)·2⁻²³ ∈ (0,1) (exact in fp32), where
    x_i = word 0 of Philox(ctr=(token id, step, uid_lo, uid_hi), key=seed)."""
    tid = np.asarray(token_ids, dtype=np.uint32)
    x = philox4x32_10((tid, step & 0xFFFFFFFF, uid & 0xFFFFFFFF, (uid >> 32) & 0xFFFFFFFF),
                      (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))[0]
    u = ((x >> np.uint32(9)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -23)
    return -np.log(u, dtype=np.float32)


def apply_repetition_penalty(logits: np.ndarray, seen: np.ndarray, penalty: float) -> np.ndarray:
    """_custom_ops.py:360-373: seen ? (l>0 ? l/p : l·p) : l  (as l·(1/p) like torch)."""
    l = logits.astype(np.float32).copy()
    if penalty != 1.0:
        p = np.float32(penalty)
        scale = np.where(l > 0, np.float32(1.0) / p, p).astype(np.float32)
        l = np.where(seen, l * scale, l).astype(np.float32)
    return l


def _rank_desc(l: np.ndarray) -> np.ndarray:
    """Indices sorted by descending logit, ties by ascending id (stable)."""
    return np.argsort(-l.astype(np.float64), kind="stable")


def kept_set(l: np.ndarray, temperature: float, min_p: float = 0.0, top_p: float = 1.0,
             top_k: int = 0, return_margin: bool = False):
    """Boolean keep-mask after temperature → min-p → top-k → top-p.

    top-p (topk_topp_sampler.py:356-364): ascending sort, softmax, cumsum, drop
    while cumsum ≤ 1−p, always keep the last  ⇔  keep token i iff the probability
    mass ranked strictly above it is < p.  ``margin`` (optional) is
    |mass_above − p| per token, so tests can skip fp-ambiguous boundary tokens."""
    l = (l.astype(np.float32) / np.float32(temperature)).astype(np.float32)
    V = l.shape[0]
    keep = np.isfinite(l) | (l == np.inf)
    m = l.max()
    if min_p > 0.0:
        e = np.exp(l - m, dtype=np.float32)
        probs = e / e.sum(dtype=np.float32)
        keep &= ~(probs < probs.max() * np.float32(min_p))
    if top_k and 0 < top_k < V:       # topk_topp_sampler.py:346-351: every tie with the k-th value stays
        lm = np.where(keep, l, -np.inf)
        keep &= lm >= np.sort(lm)[V - top_k]
    margin = np.full(V, np.inf, np.float64)
    if top_p < 1.0:
        lm = np.where(keep, l, -np.inf)
        e = np.exp(lm - m, dtype=np.float32).astype(np.float64)
        probs = e / e.sum()
        order = _rank_desc(lm)
        above = np.concatenate(([0.0], np.cumsum(probs[order])[:-1]))
        kp = np.zeros(V, bool)
        kp[order] = above < top_p
        kp[order[0]] = True
        margin[order] = np.abs(above - top_p)
        keep &= kp
    if return_margin:
        return keep, l, margin
    return keep, l


def sample_row(logits: np.ndarray, seen: np.ndarray, *, repetition_penalty: float = 1.0,
               temperature: float = 1.0, min_p: float = 0.0, top_p: float = 1.0, top_k: int = 0,
               seed: int = 0, uid: int = 0, step: int = 0) -> int:
    """One row of the sampler: returns the speech token id (no +1000 offset)."""
    l = apply_repetition_penalty(logits, seen, repetition_penalty)
    if temperature < SAMPLING_EPS:
        return int(np.argmax(l))                 # first max = lowest id on ties
    keep, lt = kept_set(l, temperature, min_p, top_p, top_k)
    ids = np.nonzero(keep)[0]
    m = lt[ids].max()
    e = np.exp(lt[ids] - m, dtype=np.float32)
    probs = e / e.sum(dtype=np.float32)
    q = exp1_variates(seed, uid, step, ids)
    return int(ids[int(np.argmax(probs / q))])


def final_probs(logits, seen, *, repetition_penalty=1.0, temperature=1.0, min_p=0.0, top_p=1.0,
                top_k=0) -> np.ndarray:
    """Distribution the race samples from (for χ² tests)."""
    l = apply_repetition_penalty(logits, seen, repetition_penalty)
    keep, lt = kept_set(l, temperature, min_p, top_p, top_k)
    e = np.where(keep, np.exp(lt - lt[keep].max(), dtype=np.float32), 0).astype(np.float64)
    return e / e.sum()
