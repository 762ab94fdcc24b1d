"""CPU oracle for the Chatterbox T3 decode path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product path
(``chatterbox_vllm_b200``) never does; it fails loudly when the CUDA library is
missing.

PARITY STATUS: the reference repo (iMash-io/chatterbox-vllm @ 8d61738) ships no tests, golden
vectors or fixtures for this path.  Pinned instead by outputs of the
reference's OWN code run in the build container (``oracle/make_golden.py``, committed under
``tests/golden/``; nothing at test time reads /root/reference):
  * **the T3 glue — pinned**: ``models/t3/t3.py`` is executed from /root/reference with its vllm
    imports stubbed; ``T3VllmModel.get_input_embeddings`` / ``split_prefill_decode`` /
    ``compute_logits``, ``T3MultiModalProcessor.apply`` and ``create_triangular_matrix`` run on
    seeded inputs (``oracle/glue_cases.py``) -> ``t3_glue.npz``; ``prefill_embeds``,
    ``decode_embed``, ``cfg_merge`` (cfg 0.0 / 0.5 / 1.3, fp32 and bf16) and ``head`` below are
    held to it bit for bit (``tests/test_reference_glue.py``);
  * the conditioning encoder — pinned the same way (``cond_enc.npz``, ``oracle/cond_ref.py``);
  * **the Llama backbone, attention and the sampler chain — unpinned by the reference** (their
    arithmetic lives in un-vendored ``vllm==0.10.0`` / ``torch 2.7.1`` kernels that cannot run
    here); pinned against what can:
  * ``transformers.LlamaModel`` (fp32) for the 30-layer backbone, llama3 RoPE
    and KV-cache decode — the same HF module the original Chatterbox wraps;
  * vLLM 0.22's own CPU-runnable pieces (llama3 ``_compute_inv_freq``,
    ``apply_top_k_top_p_pytorch``, ``apply_repetition_penalties_torch``,
    ``MinPLogitsProcessor.apply`` math) for RoPE tables and the sampler chain;
  * the reference's integer constants / tokenizer / README token totals for the
    prompt layout.

What is restated (reference file:line → function here)
  t3-model/config.json:1-33                          → T3ModelConfig
  models/t3/modules/t3_config.py:1-35                → T3ModelConfig (ids / sizes)
  models/t3/t3.py:503-514 (decode embed)             → T3Oracle.decode_embed
  models/t3/t3.py:548-569 (full-block prefill embed) → T3Oracle.prefill_embeds
  models/t3/t3.py:725-739 (CFG merge)                → cfg_merge
  models/t3/t3.py:742-778 (paired forward)           → T3Oracle.forward_rows
  vllm-0.22 model_executor/models/llama.py:212-226,316-333,394-430
                                                     → T3Oracle._layer
  vllm-0.22 layers/rotary_embedding/llama3_rope.py:33-54, base.py:59-88
                                                     → llama3_inv_freq, rope_cos_sin

Two numeric modes:
  "fp32" : weights are the bf16-rounded values up-cast; all math fp32 ("truth").
  "bf16" : same, but activations are rounded to bf16 at exactly the points where
           the reference materialises a bf16 tensor (GEMM outputs, norm outputs,
           residual stream, rotated q/k, KV cache, attention output, logits, CFG).
           This is the model the CUDA kernels are held to.

CFG semantics: "paired" (default; each of cond/uncond owns its KV history — the
north_star's intended semantics) or "none" (reference-as-executed: the uncond
rows' attention output is identically zero).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch

BF16 = torch.bfloat16


@dataclass
class T3ModelConfig:
    # t3-model/config.json (hidden_size reset to 1024, t3.py:309)
    hidden: int = 1024
    n_layers: int = 30
    n_heads: int = 16
    head_dim: int = 64
    inter: int = 4096
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_factor: float = 8.0
    rope_low: float = 1.0
    rope_high: float = 4.0
    rope_orig_max: int = 8192
    # t3_config.py:1-35
    text_vocab: int = 704
    speech_vocab: int = 8194
    start_text: int = 255
    stop_text: int = 0
    start_speech: int = 6561
    stop_speech: int = 6562
    max_text: int = 2048
    max_speech: int = 4096
    cond_len: int = 34           # CONDITIONING_SIZE, t3.py:42
    max_pos: int = 2048          # rope table length the oracle materialises

    @classmethod
    def multilingual(cls, **kw):
        return cls(text_vocab=2454, **kw)


def llama3_inv_freq(cfg: T3ModelConfig) -> torch.Tensor:
    """vllm-0.22 rotary_embedding/base.py:78-88 + llama3_rope.py:33-54 (fp32)."""
    d = cfg.head_dim
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float) / d))
    low_wl = cfg.rope_orig_max / cfg.rope_low
    high_wl = cfg.rope_orig_max / cfg.rope_high
    wl = 2 * math.pi / inv
    smooth = (cfg.rope_orig_max / wl - cfg.rope_low) / (cfg.rope_high - cfg.rope_low)
    return torch.where(
        wl < high_wl, inv,
        torch.where(wl > low_wl, inv / cfg.rope_factor,
                    (1 - smooth) * inv / cfg.rope_factor + smooth * inv))


def rope_cos_sin(cfg: T3ModelConfig, n_pos: int | None = None) -> torch.Tensor:
    """[n_pos, head_dim] fp32 = cat(cos[32], sin[32]) (base.py:90-100)."""
    n_pos = n_pos or cfg.max_pos
    t = torch.arange(n_pos, dtype=torch.float)
    fr = torch.einsum("i,j->ij", t, llama3_inv_freq(cfg))
    return torch.cat((fr.cos(), fr.sin()), dim=-1)


def _r(x: torch.Tensor, mode: str) -> torch.Tensor:
    """Round to bf16 storage in 'bf16' mode; identity in 'fp32' mode."""
    return x.to(BF16).float() if mode == "bf16" else x


@dataclass
class T3Weights:
    """bf16 weights in the layout the checkpoint implies."""
    cfg: T3ModelConfig
    wqkv: list = field(default_factory=list)     # [3H, H]  (q;k;v stacked, llama.py:433-441)
    wo: list = field(default_factory=list)       # [H, H]
    wgu: list = field(default_factory=list)      # [2I, H]  (gate;up stacked)
    wd: list = field(default_factory=list)       # [H, I]
    ln1: list = field(default_factory=list)      # [H]
    ln2: list = field(default_factory=list)      # [H]
    norm: torch.Tensor | None = None             # [H]
    speech_emb: torch.Tensor | None = None       # [V, H]
    speech_head: torch.Tensor | None = None      # [V, H]
    text_emb: torch.Tensor | None = None         # [Vt, H]
    text_pos: torch.Tensor | None = None         # [max_text+2, H]
    speech_pos: torch.Tensor | None = None       # [max_speech+4, H]

    @classmethod
    def random(cls, cfg: T3ModelConfig, seed: int = 0, head_scale: float = 1.0,
               std: float = 0.02, norm_jitter: float = 0.0) -> "T3Weights":
        """SURVEY.md §8(d) 'Weights': N(0, 0.02²) on CPU (seeded), cast to bf16,
        RMSNorm weights = 1 (+ optional jitter so the weight multiply is exercised);
        ``head_scale`` gives the 'sharp' variant with a well-conditioned argmax."""
        g = torch.Generator().manual_seed(seed)

        def n(*shape, s=std):
            return (torch.randn(*shape, generator=g) * s).to(BF16)

        def ln():
            w = torch.ones(cfg.hidden)
            if norm_jitter:
                w = w + norm_jitter * torch.randn(cfg.hidden, generator=g)
            return w.to(BF16)

        H, I = cfg.hidden, cfg.inter
        w = cls(cfg)
        for _ in range(cfg.n_layers):
            w.wqkv.append(n(3 * H, H))
            w.wo.append(n(H, H))
            w.wgu.append(n(2 * I, H))
            w.wd.append(n(H, I))
            w.ln1.append(ln())
            w.ln2.append(ln())
        w.norm = ln()
        w.speech_emb = n(cfg.speech_vocab, H)
        w.speech_head = n(cfg.speech_vocab, H, s=std * head_scale)
        w.text_emb = n(cfg.text_vocab, H)
        w.text_pos = n(cfg.max_text + 2, H)
        w.speech_pos = n(cfg.max_speech + 4, H)
        return w

    def state_dict(self) -> dict:
        """Checkpoint key layout the reference loader consumes (t3.py:347-369,
        SURVEY.md App. A): un-stacked q/k/v and gate/up."""
        H, I = self.cfg.hidden, self.cfg.inter
        sd = {}
        for i in range(self.cfg.n_layers):
            p = f"tfmr.layers.{i}."
            q, k, v = self.wqkv[i].split([H, H, H], dim=0)
            g, u = self.wgu[i].split([I, I], dim=0)
            sd[p + "self_attn.q_proj.weight"] = q
            sd[p + "self_attn.k_proj.weight"] = k
            sd[p + "self_attn.v_proj.weight"] = v
            sd[p + "self_attn.o_proj.weight"] = self.wo[i]
            sd[p + "mlp.gate_proj.weight"] = g
            sd[p + "mlp.up_proj.weight"] = u
            sd[p + "mlp.down_proj.weight"] = self.wd[i]
            sd[p + "input_layernorm.weight"] = self.ln1[i]
            sd[p + "post_attention_layernorm.weight"] = self.ln2[i]
        sd["tfmr.norm.weight"] = self.norm
        sd["speech_emb.weight"] = self.speech_emb
        sd["speech_head.weight"] = self.speech_head
        sd["text_emb.weight"] = self.text_emb
        sd["text_pos_emb.emb.weight"] = self.text_pos
        sd["speech_pos_emb.emb.weight"] = self.speech_pos
        return sd


def fp8_e4m3_rows(w: torch.Tensor):
    """Per-output-feature FP8-e4m3 quantisation of an nn.Linear weight [N, K] (the optional weight-only
    FP8 path, SURVEY.md §8(d) cfg 5): scale[n] = max|w[n,:]| / 448 (1 for a zero row),
    q = e4m3_rn(w / scale).  Returns (q as torch.float8_e4m3fn, scale fp32)."""
    wf = w.float()
    amax = wf.abs().amax(dim=1)
    scale = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
    return (wf / scale[:, None]).to(torch.float8_e4m3fn), scale


def with_fp8_projections(w: "T3Weights") -> "T3Weights":
    """The weights the FP8 engine effectively multiplies by: q·scale in fp32 for the four projections
    of every layer and the speech head (embeddings, norms, position tables stay bf16).  T3Oracle
    upcasts weights with .float(), so it runs on these unchanged."""
    import copy
    dq = lambda t: (lambda q, s: q.float() * s[:, None])(*fp8_e4m3_rows(t))  # noqa: E731
    w8 = copy.copy(w)
    w8.wqkv, w8.wo = [dq(t) for t in w.wqkv], [dq(t) for t in w.wo]
    w8.wgu, w8.wd = [dq(t) for t in w.wgu], [dq(t) for t in w.wd]
    w8.speech_head = dq(w.speech_head)
    return w8


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float, mode: str) -> torch.Tensor:
    """vLLM rms_norm: out = bf16(bf16(x·rstd) · w); fp32 statistics."""
    xf = x.float()
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return _r(_r(xf * rstd, mode) * w.float(), mode)


def apply_rope(x: torch.Tensor, pos: torch.Tensor, cs: torch.Tensor, mode: str) -> torch.Tensor:
    """neox-style rotation of [T, nh, d] (rotary_embedding/common.py:167-168);
    cos/sin are read from a model-dtype cache (base.py:59-61)."""
    half = x.shape[-1] // 2
    c = _r(cs[pos, :half], mode)[:, None, :]
    s = _r(cs[pos, half:], mode)[:, None, :]
    x1, x2 = x[..., :half].float(), x[..., half:].float()
    return _r(torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), dim=-1), mode)


def cfg_merge(cond: torch.Tensor, uncond: torch.Tensor, scale: float, mode: str) -> torch.Tensor:
    """t3.py:732  logits = cond + cfg·(cond − uncond), evaluated in the logits
    dtype (bf16 in the reference: one rounding per torch op)."""
    d = _r(cond - uncond, mode)
    return _r(cond + _r(scale * d, mode), mode)


class SeqState:
    """Dense per-sequence KV history (paging is bookkeeping only — checked
    separately, bit-exact, in paging_ref.py)."""

    def __init__(self, cfg: T3ModelConfig):
        self.k = [torch.zeros(0, cfg.n_heads, cfg.head_dim) for _ in range(cfg.n_layers)]
        self.v = [torch.zeros(0, cfg.n_heads, cfg.head_dim) for _ in range(cfg.n_layers)]
        self.len = 0


class T3Oracle:
    def __init__(self, weights: T3Weights, mode: str = "bf16", uncond_attention: str = "paired"):
        assert mode in ("fp32", "bf16") and uncond_attention in ("paired", "none")
        self.w, self.cfg, self.mode = weights, weights.cfg, mode
        self.uncond_attention = uncond_attention
        self.cs = rope_cos_sin(self.cfg)
        self.trace: dict | None = None   # set to {} to record per-layer hidden states

    # ---- embeddings -----------------------------------------------------
    def decode_embed(self, speech_ids: torch.Tensor) -> torch.Tensor:
        """t3.py:503-514: speech_emb[id] for BOTH pair members; no positional
        embedding during decode."""
        return self.w.speech_emb[speech_ids].float()

    def prefill_embeds(self, cond_emb: torch.Tensor, text_ids: torch.Tensor):
        """t3.py:548-569: cond = cond34 ‖ text_emb+text_pos ‖ SOS+speech_pos[0];
        uncond = same with the text rows zeroed.  Sums are taken in model dtype."""
        m = self.mode
        n = len(text_ids)
        text = _r(self.w.text_emb[text_ids].float() + self.w.text_pos[:n].float(), m)
        sos = _r(self.w.speech_emb[self.cfg.start_speech].float() + self.w.speech_pos[0].float(), m)[None]
        c34 = _r(cond_emb.float(), m)
        cond = torch.cat((c34, text, sos), 0)
        uncond = torch.cat((c34, torch.zeros_like(text), sos), 0)
        return cond, uncond

    # ---- backbone -------------------------------------------------------
    def _layer(self, li, x, res, rows, zero_attn):
        """One decoder layer over T rows.  rows = [(SeqState, n_new)] in row order.
        vllm-0.22 llama.py:316-333 (residual convention) and :212-226 (attention)."""
        cfg, w, m = self.cfg, self.w, self.mode
        H, nh, d = cfg.hidden, cfg.n_heads, cfg.head_dim
        if res is None:
            res = x
        else:
            res = _r(x + res, m)
        h = rmsnorm(res, w.ln1[li], cfg.rms_eps, m)
        qkv = _r(h @ w.wqkv[li].float().T, m)
        q, k, v = qkv.split([H, H, H], dim=-1)
        T = x.shape[0]
        pos = torch.cat([torch.arange(st.len, st.len + n) for st, n in rows])
        q = apply_rope(q.view(T, nh, d), pos, self.cs, m)
        k = apply_rope(k.view(T, nh, d), pos, self.cs, m)
        v = v.reshape(T, nh, d)
        attn = torch.zeros(T, nh, d)
        r0 = 0
        for ri, (st, n) in enumerate(rows):
            st.k[li] = torch.cat((st.k[li], k[r0:r0 + n]), 0)
            st.v[li] = torch.cat((st.v[li], v[r0:r0 + n]), 0)
            if not zero_attn[ri]:
                L0 = st.len
                K, V = st.k[li], st.v[li]                       # [L0+n, nh, d]
                s = torch.einsum("thd,lhd->htl", q[r0:r0 + n], K) * (d ** -0.5)
                causal = torch.arange(L0 + n)[None, :] <= (L0 + torch.arange(n))[:, None]
                s = s.masked_fill(~causal[None], float("-inf"))
                p = torch.softmax(s, dim=-1)
                attn[r0:r0 + n] = torch.einsum("htl,lhd->thd", p, V)
            r0 += n
        attn = _r(attn.reshape(T, H), m)
        o = _r(attn @ w.wo[li].float().T, m)
        res = _r(o + res, m)
        h = rmsnorm(res, w.ln2[li], cfg.rms_eps, m)
        gu = _r(h @ w.wgu[li].float().T, m)
        g, u = gu.split([cfg.inter, cfg.inter], dim=-1)
        act = _r(torch.nn.functional.silu(g) * u, m)
        dn = _r(act @ w.wd[li].float().T, m)
        if self.trace is not None:
            self.trace.setdefault("layers", []).append((dn.clone(), res.clone()))
        return dn, res

    def forward_rows(self, embeds, rows, zero_attn=None):
        """embeds [T, H] (already in the numeric mode's precision); rows as in
        _layer.  Returns final-norm hidden states [T, H] and advances each
        SeqState.len."""
        zero_attn = zero_attn or [False] * len(rows)
        x, res = embeds, None
        for li in range(self.cfg.n_layers):
            x, res = self._layer(li, x, res, rows, zero_attn)
        res = _r(x + res, self.mode)
        out = rmsnorm(res, self.w.norm, self.cfg.rms_eps, self.mode)
        for st, n in rows:
            st.len += n
        return out

    def head(self, h: torch.Tensor) -> torch.Tensor:
        return _r(h @ self.w.speech_head.float().T, self.mode)

    # ---- CFG pair API ---------------------------------------------------
    def new_pair(self):
        return SeqState(self.cfg), SeqState(self.cfg)

    def prefill_pair(self, pair, cond_emb, text_ids, cfg_scale):
        """Prefill one CFG pair's whole prefix in one pass; returns merged logits
        [V] for the first speech token."""
        c, u = self.prefill_embeds(cond_emb, torch.as_tensor(text_ids))
        P = c.shape[0]
        none = self.uncond_attention == "none"
        h = self.forward_rows(torch.cat((c, u), 0), [(pair[0], P), (pair[1], P)], [False, none])
        lg = self.head(torch.stack((h[P - 1], h[2 * P - 1])))
        return cfg_merge(lg[0], lg[1], cfg_scale, self.mode)

    def decode_pairs(self, pairs, prev_tokens, cfg_scale):
        """One decode step for a batch of pairs (rows: all cond, then all uncond —
        t3.py:768-773).  prev_tokens are speech ids (no +1000 offset).  Returns
        merged logits [B, V]."""
        B = len(pairs)
        e = _r(self.decode_embed(torch.as_tensor(prev_tokens)), self.mode)
        rows = [(p[0], 1) for p in pairs] + [(p[1], 1) for p in pairs]
        none = self.uncond_attention == "none"
        h = self.forward_rows(torch.cat((e, e), 0), rows, [False] * B + [none] * B)
        lg = self.head(h)
        return cfg_merge(lg[:B], lg[B:], cfg_scale, self.mode)
