"""T3Engine — host side of the B200 T3 speech-token path.

Replaces, for this path only, what the reference obtains from ``vllm.LLM(...)`` +
``T3VllmModel`` (tts.py:130-140,408-418; models/t3/t3.py:302-781): weight loading with q/k/v
and gate/up stacking (t3.py:347-379), prompt layout (t3.py:229-262,548-569), paged-KV
bookkeeping (vLLM block pool / block table), the paired cond/uncond forward and the sampler.
All arithmetic runs in libcbx_t3.so; this file only moves pointers and small index arrays."""
from __future__ import annotations

import ctypes as C
import math
import os
from dataclasses import dataclass, field

import numpy as np
import torch

from . import ops
from ._lib import (BLOCK_SIZE, EngineCfg, Fp8Weights, HIDDEN, PAGE_BYTES, SPEECH_VOCAB, Weights, check, load)

BF16 = torch.bfloat16
COND_LEN = 34                 # CONDITIONING_SIZE, t3.py:42
START_SPEECH, STOP_SPEECH = 6561, 6562
SPEECH_TOKEN_OFFSET = 1000    # t3.py:49


def llama3_cos_sin(max_pos: int, head_dim: int = 64, theta: float = 500000.0, factor: float = 8.0,
                   low: float = 1.0, high: float = 4.0, orig_max: int = 8192) -> torch.Tensor:
    """bf16 [max_pos, 64] = cos(32) ‖ sin(32); llama3 rescale as t3-model/config.json:19-26
    (vllm rotary_embedding/llama3_rope.py:33-54, base.py:90-100)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
    wl = 2 * math.pi / inv
    smooth = (orig_max / wl - low) / (high - low)
    inv = torch.where(wl < orig_max / high, inv,
                      torch.where(wl > orig_max / low, inv / factor, (1 - smooth) * inv / factor + smooth * inv))
    fr = torch.einsum("i,j->ij", torch.arange(max_pos, dtype=torch.float), inv)
    return torch.cat((fr.cos(), fr.sin()), dim=-1).to(BF16)


@dataclass
class SamplingParams:
    """The vLLM SamplingParams fields the reference actually uses (tts.py:408-418)."""
    temperature: float = 0.8
    top_p: float = 1.0
    min_p: float = 0.0
    top_k: int = 0
    repetition_penalty: float = 1.0
    max_tokens: int = 1000
    seed: int | None = None
    ignore_eos: bool = False
    stop_token_ids: list = field(default_factory=lambda: [STOP_SPEECH + SPEECH_TOKEN_OFFSET])


@dataclass
class Request:
    text_ids: list            # tokenizer output for "[START]…[STOP]" (n_text ids)
    cond_emb: torch.Tensor    # [34, 1024] fp32 (CPU or CUDA)
    params: SamplingParams
    uid: int = 0
    penalize_prompt_ids: bool = False   # MTL quirk: prompt ids ≥1000 hit speech id−1000
    history: list = field(default_factory=list)   # scheduler-internal: speech ids (no offset) this request had already
                                                  # emitted when it was preempted; its KV is recomputed over them


class T3Engine:
    def __init__(self, state_dict: dict, *, device="cuda:0", max_pairs: int = 128, max_model_len: int = 1000,
                 num_blocks: int | None = None, max_rows: int | None = None, cfg_scale: float | None = None,
                 rms_eps: float = 1e-5, base_seed: int = 0, use_graph: bool = True,
                 stop_token: int = STOP_SPEECH, weight_dtype: str | None = None):
        """weight_dtype: "bf16" (default; env CBX_WEIGHT_DTYPE overrides) or "fp8_e4m3" — the optional
        weight-only FP8 path: the four projections of every layer and the speech head are quantised per
        output feature on the GPU at load time and the bf16 copies are dropped."""
        load()                                           # fail loudly if the CUDA library is missing
        if not torch.cuda.is_available():
            raise RuntimeError("T3Engine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.max_pairs, self.max_model_len = max_pairs, max_model_len
        # CHATTERBOX_CFG_SCALE is read once at model construction (t3.py:343)
        self.cfg_scale = float(os.environ.get("CHATTERBOX_CFG_SCALE", "0.5")) if cfg_scale is None else cfg_scale
        self.use_graph, self.base_seed, self._next_uid = use_graph, base_seed, 0
        # a dedicated non-default stream: CUDA graphs cannot be captured on the legacy stream
        self.stream = torch.cuda.Stream(device=self.device)
        self._copy_stream = torch.cuda.Stream(device=self.device)   # host->device copies of prefill metadata (prefill())
        self._load_weights(state_dict, rms_eps)
        self.bt_stride = -(-max_model_len // BLOCK_SIZE)
        if num_blocks is None:
            num_blocks = 2 * max_pairs * self.bt_stride + 1
        self.num_blocks = num_blocks
        # step-buffer rows = the largest prefill chunk.  17 408 rows = 136 row tiles: the prefill GEMM grids then come to whole
        # numbers of waves on 148 SMs x 2 CTAs (measured at M1: three chunks of 17 067 rows 65.6 ms, four of 12 800 67.9 ms)
        self.max_rows = max_rows or max(2 * max_pairs, 17408)
        self.kv_cache = torch.zeros(self.n_layers * num_blocks * PAGE_BYTES, dtype=torch.uint8, device=self.device)
        self.block_table = torch.zeros(2 * max_pairs, self.bt_stride, dtype=torch.int32, device=self.device)
        self.pool = ops.BlockPool(num_blocks)
        self.sb = ops.SamplerBuffers(max_pairs, max_model_len, self.device, stop_token=stop_token,
                                     cfg_scale=self.cfg_scale)
        self._cfg = EngineCfg(max_pairs, self.max_rows, max_model_len, num_blocks, max_model_len)
        nbytes = load().cbx_engine_workspace_bytes(C.byref(self._cfg))
        self.workspace = torch.zeros(nbytes + 1024, dtype=torch.uint8, device=self.device)
        ws = (self.workspace.data_ptr() + 1023) // 1024 * 1024
        self._st = self.sb.struct()
        self._h = load().cbx_engine_create(C.byref(self._cfg), C.byref(self._w), self.kv_cache.data_ptr(),
                                           self.block_table.data_ptr(), ws, C.byref(self._st))
        if not self._h:
            raise RuntimeError("cbx_engine_create: " + load().cbx_last_error().decode())
        self.weight_dtype = (weight_dtype or os.environ.get("CBX_WEIGHT_DTYPE", "bf16")).lower()
        if self.weight_dtype in ("fp8", "fp8_e4m3", "e4m3"):
            self.weight_dtype = "fp8_e4m3"
            self._quantize_weights()
        elif self.weight_dtype != "bf16":
            raise ValueError(f"weight_dtype {self.weight_dtype!r}: expected 'bf16' or 'fp8_e4m3'")
        elif os.environ.get("CBX_MLP_FUSED") == "1":
            # opt-in two-launch layer of the T <= 4 decode path (measured equal to the default three launches): its fused
            # MLP block reads the down-proj weights by MLP feature, so a transposed copy (+8 MB per layer) keeps a CTA's
            # slice contiguous (csrc/k_mlp.cu)
            self.wd_t = [w.t().contiguous() for w in self.wd]
            self._arr_wd_t = (C.c_void_p * self.n_layers)(*[t.data_ptr() for t in self.wd_t])
            torch.cuda.synchronize()
            check(load().cbx_engine_set_down_t(self._h, self._arr_wd_t, self.n_layers), "set_down_t")
        self.gpu_launches = 0
        self._rows: dict[int, list[int]] = {}             # block-table row -> its blocks, in table order
        self.preemptions = 0

    # ------------------------------------------------------------------ weights
    def _load_weights(self, sd: dict, rms_eps: float) -> None:
        """Checkpoint key layout of t3_cfg.safetensors; q/k/v and gate/up are
        stacked as vLLM's loader does (vllm llama.py:433-441); gate/up then packed for the
        fused SiLU·mul epilogue."""
        dev = self.device
        g = lambda k: sd[k].to(device=dev, dtype=BF16).contiguous()  # noqa: E731
        L = 0
        while f"tfmr.layers.{L}.self_attn.q_proj.weight" in sd:
            L += 1
        if L == 0:
            raise KeyError("state_dict has no tfmr.layers.* keys")
        self.n_layers = L
        self.wqkv, self.wo, self.wgu, self.wd, self.ln1, self.ln2 = [], [], [], [], [], []
        for i in range(L):
            p = f"tfmr.layers.{i}."
            self.wqkv.append(torch.cat([g(p + f"self_attn.{n}_proj.weight") for n in "qkv"], 0).contiguous())
            self.wo.append(g(p + "self_attn.o_proj.weight"))
            gu = torch.cat([g(p + "mlp.gate_proj.weight"), g(p + "mlp.up_proj.weight")], 0).contiguous()
            self.wgu.append(ops.pack_gate_up(gu))
            self.wd.append(g(p + "mlp.down_proj.weight"))
            self.ln1.append(g(p + "input_layernorm.weight"))
            self.ln2.append(g(p + "post_attention_layernorm.weight"))
        self.norm = g("tfmr.norm.weight")
        self.speech_emb, self.speech_head = g("speech_emb.weight"), g("speech_head.weight")
        self.text_emb, self.text_pos = g("text_emb.weight"), g("text_pos_emb.emb.weight")
        self.speech_pos = g("speech_pos_emb.emb.weight")
        assert self.speech_emb.shape == (SPEECH_VOCAB, HIDDEN)
        self.cos_sin = llama3_cos_sin(max(self.max_model_len + 8, 2048)).to(dev)
        arr = lambda ts: (C.c_void_p * L)(*[t.data_ptr() for t in ts])  # noqa: E731
        self._arrs = [arr(t) for t in (self.wqkv, self.wo, self.wgu, self.wd, self.ln1, self.ln2)]
        self._w = Weights(L, *self._arrs, self.norm.data_ptr(), self.speech_emb.data_ptr(),
                          self.speech_head.data_ptr(), self.cos_sin.data_ptr(), rms_eps)
        torch.cuda.synchronize()

    def _quantize_weights(self) -> None:
        """bf16 projections -> e4m3 bytes + per-output-feature fp32 scales (gate/up AFTER packing, so the
        scale array is in packed row order); the engine is switched over and the bf16 copies released."""
        L = self.n_layers
        with torch.cuda.stream(self.stream):
            q = lambda ws: [ops.quantize_fp8_rows(w) for w in ws]  # noqa: E731
            self.wqkv, self.wo, self.wgu, self.wd = q(self.wqkv), q(self.wo), q(self.wgu), q(self.wd)
            self.speech_head = ops.quantize_fp8_rows(self.speech_head)
        self.stream.synchronize()
        arr = lambda ps: (C.c_void_p * L)(*ps)  # noqa: E731
        self._arrs8 = [arr([w.q.data_ptr() for w in ws]) for ws in (self.wqkv, self.wo, self.wgu, self.wd)] + \
                      [arr([w.scale_buf.data_ptr() for w in ws]) for ws in (self.wqkv, self.wo, self.wgu, self.wd)]
        self._w8 = Fp8Weights(L, *self._arrs8, self.speech_head.q.data_ptr(), self.speech_head.scale_buf.data_ptr())
        check(load().cbx_engine_set_fp8_weights(self._h, C.byref(self._w8)), "set_fp8_weights")

    def projection_weight_bytes(self) -> int:
        """Bytes one decode step streams for the projections + head (the W term of the step roofline)."""
        ws = self.wqkv + self.wo + self.wgu + self.wd + [self.speech_head]
        if self.weight_dtype == "fp8_e4m3":
            return sum(w.q.numel() + 4 * w.q.shape[0] for w in ws)
        return sum(w.numel() * 2 for w in ws)

    # ------------------------------------------------------------------ helpers
    def buffer(self, name: str, shape, dtype) -> torch.Tensor:
        """Zero-copy torch view of an engine-internal device buffer."""
        ptr = load().cbx_engine_buffer(self._h, name.encode())
        if not ptr:
            raise KeyError(name)
        n = int(np.prod(shape))
        esz = torch.empty(0, dtype=dtype).element_size()
        off = ptr - self.workspace.data_ptr()
        return self.workspace[off:off + n * esz].view(dtype).view(*shape)

    def _stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream

    def set_trace(self, trace: torch.Tensor | None) -> None:
        """Parity hook: bf16 [n_layers, T, 1024] receives the residual stream after every layer."""
        check(load().cbx_engine_set_trace(self._h, trace.data_ptr() if trace is not None else None), "set_trace")

    def forward(self, T: int, max_kv_len: int) -> None:
        check(load().cbx_engine_forward(self._h, T, max_kv_len, self._stream()), "engine_forward")

    def sample(self, B: int) -> None:
        check(load().cbx_engine_sample(self._h, B, self._stream()), "engine_sample")

    def decode(self, B: int, n_steps: int, max_kv_len: int) -> None:
        check(load().cbx_engine_decode(self._h, B, n_steps, max_kv_len, int(self.use_graph), self._stream()),
              "engine_decode")
        self.gpu_launches += load().cbx_engine_last_launch_count(self._h)

    # ------------------------------------------------------------------ admission + prefill
    def _reset_slots(self, B: int) -> None:
        """Start a generate() call with B empty pair slots: every slot finished, staged inputs inert
        (kv_len 0 => attention skips the row, slot -1 => nothing is cached)."""
        self._B = B
        sb = self.sb
        sb.finished[:B] = 1
        sb.n_out[:B] = 0
        sb.n_finished.zero_()
        sb.seq_len[:2 * B] = 0
        self.buffer("pos", (2 * B,), torch.int32).zero_()
        self.buffer("kv_len", (2 * B,), torch.int32).zero_()
        self.buffer("slot", (2 * B,), torch.int64).fill_(-1)
        self.buffer("row_seq", (2 * B,), torch.int32).copy_(torch.arange(2 * B, dtype=torch.int32, device=self.device))

    def _admit(self, reqs: list[Request], slots: list[int] | None = None,
               lookahead: int | None = None) -> tuple[list[int], list[list[int]]]:
        """Give both members of every pair (rows b and B+b of the block table, B = slot count of this call) their
        first KV blocks and upload those rows.  lookahead=None reserves prompt + max_tokens up front (callers that
        drive decode() themselves: bench, tests); lookahead=k allocates as vLLM does — cdiv(tokens, 16) blocks for
        the tokens that can exist before the next top-up (prefix + first token + k), the rest arriving through
        _grow() (vllm v1/core/single_type_kv_cache_manager.py:120).  With slots=None the requests take slots
        0..n-1 of a fresh n-slot batch.  Returns prefix lengths and the rows' block lists (the lists grow in place)."""
        if slots is None:
            slots = list(range(len(reqs)))
            self._reset_slots(len(reqs))
        B = self._B
        P, held, rows, ids = [], [], [], []
        for r in reqs:
            p = COND_LEN + len(r.text_ids) + 1
            if p + r.params.max_tokens > self.max_model_len:
                raise ValueError(f"prompt ({p}) + max_tokens ({r.params.max_tokens}) exceeds max_model_len")
            P.append(p + len(r.history))                   # rows to prefill: the prefix + what was generated before a preemption
        for half in range(2):
            for j, r in enumerate(reqs):
                n_tok = P[j] - len(r.history) + r.params.max_tokens
                if lookahead is not None:
                    n_tok = min(n_tok, P[j] + 1 + lookahead)
                blks = self.pool.alloc(-(-n_tok // BLOCK_SIZE))
                row = np.zeros(self.bt_stride, np.int32)
                row[:len(blks)] = blks
                rows.append(row)
                ids.append(half * B + slots[j])
                held.append(blks)
                self._rows[half * B + slots[j]] = blks
        self.block_table.index_copy_(0, torch.tensor(ids, device=self.device),
                                     torch.from_numpy(np.stack(rows)).to(self.device))
        return P, held

    def _grow(self, want: list[tuple[int, int]]) -> bool:
        """Top block-table rows up so that row r can hold n tokens, for (r, n) in `want` — in that order, cdiv(n, 16)
        blocks per row, new blocks popped from the pool's free queue exactly as vLLM's KV-cache manager does while a
        request decodes.  All or nothing: returns False (and allocates nothing) when the pool cannot serve every row."""
        need = [(r, -(-n // BLOCK_SIZE) - len(self._rows[r])) for r, n in want]
        if sum(k for _, k in need if k > 0) > self.pool.num_free():
            return False
        rr, cc, bb = [], [], []
        for r, k in need:
            if k <= 0:
                continue
            blks = self.pool.alloc(k)
            n0 = len(self._rows[r])
            rr += [r] * k
            cc += list(range(n0, n0 + k))
            bb += blks
            self._rows[r].extend(blks)
        if rr:
            dev = self.device
            self.block_table.index_put_((torch.tensor(rr, device=dev), torch.tensor(cc, device=dev)),
                                        torch.tensor(bb, dtype=torch.int32, device=dev))
        return True

    def _release(self, held: list[list[int]]) -> None:
        for blks in held:                                  # vLLM frees a request's blocks in reverse order
            self.pool.free(list(reversed(blks)))
            for r in [r for r, v in self._rows.items() if v is blks]:
                del self._rows[r]

    def _park_slot(self, b: int) -> None:
        """Make pair slot b inert on the device (as _reset_slots leaves an empty slot): finished, nothing staged —
        kv_len 0 so attention skips its rows, slot -1 so nothing is written to blocks it no longer owns."""
        B = self._B
        idx = torch.tensor([b, B + b], device=self.device)
        self.sb.finished[b] = 1
        self.buffer("kv_len", (2 * B,), torch.int32)[idx] = 0
        self.buffer("slot", (2 * B,), torch.int64)[idx] = -1

    def _init_sampler_state(self, reqs: list[Request], P: list[int], slots: list[int] | None = None) -> None:
        sb, dev, B = self.sb, self.device, self._B
        slots = list(range(len(reqs))) if slots is None else slots
        idx = torch.tensor(slots, device=dev)
        up = lambda vals, dt: torch.tensor(vals, dtype=dt).to(dev)  # noqa: E731
        pr = [r.params for r in reqs]
        sb.temperature[idx] = up([p.temperature for p in pr], torch.float32)
        sb.top_p[idx] = up([p.top_p for p in pr], torch.float32)
        sb.min_p[idx] = up([p.min_p for p in pr], torch.float32)
        sb.rep_pen[idx] = up([p.repetition_penalty for p in pr], torch.float32)
        sb.top_k[idx] = up([p.top_k for p in pr], torch.int32)
        sb.ignore_eos[idx] = up([int(p.ignore_eos) for p in pr], torch.int32)
        sb.max_out[idx] = up([p.max_tokens for p in pr], torch.int32)
        seeds = [(p.seed if p.seed is not None else self.base_seed) & (2**63 - 1) for p in pr]
        sb.seed[idx] = up(seeds, torch.int64)
        sb.uid[idx] = up([r.uid for r in reqs], torch.int64)
        sb.n_out[idx] = up([len(r.history) for r in reqs], torch.int32)
        for j, r in enumerate(reqs):                       # a recomputed request keeps the tokens it had emitted
            if r.history:
                sb.out_tokens[slots[j], :len(r.history)] = up(r.history, torch.int32)
        sb.finished[idx] = 0
        # cache holds P tokens after prefill; the prefill-time sample call adds 1, so start at P-1
        lens = up([p - 1 for p in P], torch.int32)
        sb.seq_len[idx] = lens
        sb.seq_len[idx + B] = lens
        seen = np.zeros((len(reqs), ops.SEEN_WORDS), np.uint32)
        for j, r in enumerate(reqs):
            if r.penalize_prompt_ids:
                for t in r.text_ids:
                    if t >= SPEECH_TOKEN_OFFSET and t - SPEECH_TOKEN_OFFSET < SPEECH_VOCAB:
                        sp_id = t - SPEECH_TOKEN_OFFSET
                        seen[j, sp_id >> 5] |= np.uint32(1 << (sp_id & 31))
            for sp_id in r.history:
                seen[j, sp_id >> 5] |= np.uint32(1 << (sp_id & 31))
        sb.seen[idx] = torch.from_numpy(seen.view(np.int32)).to(dev)

    def prefill(self, reqs: list[Request], P: list[int], slots: list[int] | None = None) -> None:
        """One-pass prefill of every new pair's whole prefix (cond rows then uncond rows per pair),
        chunked by max_rows; then the first speech token of every new pair is sampled.  When other
        slots of the batch are mid-decode, their staged next-step inputs are preserved."""
        n, dev, B = len(reqs), self.device, self._B
        slots = list(range(n)) if slots is None else slots
        whole = n == B
        staged = None
        if not whole:                                        # prefill reuses the step buffers: save the running pairs' rows
            names = (("res", (2 * B, HIDDEN), BF16), ("pos", (2 * B,), torch.int32), ("slot", (2 * B,), torch.int64),
                     ("kv_len", (2 * B,), torch.int32), ("row_seq", (2 * B,), torch.int32))
            staged = [(self.buffer(*nm), self.buffer(*nm).clone()) for nm in names]
        # the conditioning rows go up request by request (pinned host tensors copy asynchronously; a torch.stack on the host
        # would make one pageable 18 MB tensor at M1 and a blocking copy)
        cond = torch.empty(n, COND_LEN, HIDDEN, dtype=torch.float32, device=dev)
        for i, r in enumerate(reqs):
            cond[i].copy_(r.cond_emb, non_blocking=True)
        h_last = torch.empty(2 * n, HIDDEN, dtype=BF16, device=dev)
        x = self.buffer("x", (self.max_rows, HIDDEN), BF16)
        h = self.buffer("h", (self.max_rows, HIDDEN), BF16)
        # ---- per chunk: host work, host->device copies on a COPY stream, then the chunk's kernels on the engine's stream.
        # (With the copies on the engine's stream each pageable copy waited, in stream order, for the previous chunk's
        # 16 ms of GPU work, and the host work of the next chunk — ~2 ms — then sat exposed between chunks: 8 ms of a 74 ms
        # M1 prefill.  Now the host plans chunk k+1 while the GPU runs chunk k.)
        # Chunks of at most max_rows rows, of EQUAL size: a short last chunk (M1: 40 + 40 + 40 + 8 pairs) runs the small-M
        # GEMM configurations at a fraction of the big chunks' rate.  A pair joins chunk k while its midpoint lies in front
        # of the k-th equal cut of the row total.
        main = torch.cuda.current_stream()
        def up(v, dtype=torch.int32):                         # noqa: E306  host list -> device tensor, on the copy stream
            with torch.cuda.stream(self._copy_stream):
                t = torch.tensor(v if len(v) else [0], dtype=dtype).to(dev)
            t.record_stream(main)
            return t
        i32 = up
        total_rows = 2 * sum(P)
        target = total_rows / max(1, -(-total_rows // self.max_rows))
        plan = []
        j0 = rows_done = k_chunk = 0
        while j0 < n:
            j1, rows = j0, 0
            k_chunk += 1
            while j1 < n and rows + 2 * P[j1] <= self.max_rows and (rows == 0 or rows_done + rows + P[j1] <= k_chunk * target):
                rows += 2 * P[j1]
                j1 += 1
            rows_done += rows
            if j1 == j0:
                raise ValueError("max_rows too small for one pair's prefix")
            # one record per SEQUENCE goes to the device; the per-row descriptors, positions, block-table rows and context
            # lengths (51 200 rows at M1) are written there, straight into the step buffers (cbx_prefill_rows)
            seq, text_flat, hist_flat, last = [], [], [], {}
            starts, plens, btrows = [], [], []
            T = t_off = h_off = 0
            for j in range(j0, j1):
                nt, nh = len(reqs[j].text_ids), len(reqs[j].history)
                text_flat.extend(int(t) for t in reqs[j].text_ids)
                hist_flat.extend(int(t) for t in reqs[j].history)   # generated tokens: decode embeddings, both members
                for half in range(2):
                    row = half * B + slots[j]
                    seq.append((T, P[j], nt, j, half, row, t_off, h_off))
                    starts.append(T), plens.append(P[j]), btrows.append(row)
                    T += P[j]
                    last[half * n + j] = T - 1
                t_off += nt
                h_off += nh
            items_h = ops.prefill_items(starts, plens, btrows)             # causal attention: 16 query rows per item
            if os.environ.get("CBX_PREFILL_ROWWISE") == "1":      # debugging aid: causal prefill row by row through the decode kernel
                items_h = items_h[:0]
            with torch.cuda.stream(self._copy_stream):
                items = items_h.to(dev)
            items.record_stream(main)
            seqs = sorted(last)
            max_p, d_seq, d_text, d_hist = max(P[j0:j1]), i32(seq).view(-1, 8), i32(text_flat), i32(hist_flat)
            dst_rows, src_rows = up(seqs, torch.int64), up([last[q] for q in seqs], torch.int64)
            main.wait_stream(self._copy_stream)
            plan.append((d_seq, d_text, d_hist, items, dst_rows, src_rows))   # alive until the chunks have been enqueued
            j0 = j1
            # ---- the chunk's kernels
            d_pos, d_rseq = self.buffer("pos", (T,), torch.int32), self.buffer("row_seq", (T,), torch.int32)
            d_desc, _, _, _ = ops.prefill_rows(d_seq, d_text, d_hist, T, pos=d_pos, row_seq=d_rseq,
                                               kv_len=self.buffer("kv_len", (T,), torch.int32))
            self.buffer("slot", (T,), torch.int64).copy_(ops.slot_mapping(self.block_table, d_rseq, d_pos))
            ops.prefill_embed(d_desc, cond, self.text_emb, self.text_pos, self.speech_emb, self.speech_pos, out=x[:T])
            check(load().cbx_engine_set_prefill_items(self._h, items.data_ptr() if items.shape[0] else None, items.shape[0]),
                  "set_prefill_items")
            try:
                self.forward(T, max_p)
            finally:
                check(load().cbx_engine_set_prefill_items(self._h, None, 0), "set_prefill_items")
            h_last[dst_rows] = h[src_rows]
        h[:2 * n].copy_(h_last)
        self.buffer("head_rows", (2 * n,), torch.int32).copy_(torch.arange(2 * n, dtype=torch.int32, device=dev))
        if whole:
            self.buffer("row_seq", (2 * B,), torch.int32).copy_(torch.arange(2 * B, dtype=torch.int32, device=dev))
            self.sample(B)
            return
        # restore the running pairs' staged rows, then sample only the new pairs (which stages theirs)
        for buf, saved in staged:
            buf[: 2 * B] = saved                               # prefill clobbered rows [0, T): put everything back …
        pair_map = torch.tensor(slots, dtype=torch.int32, device=dev)
        check(load().cbx_engine_sample_subset(self._h, B, n, pair_map.data_ptr(), self._stream()), "engine_sample_subset")

    # ------------------------------------------------------------------ public entry
    def generate(self, reqs: list[Request], poll_every: int = 32) -> list[list[int]]:
        """Runs the requests through min(max_pairs, n) pair slots with continuous admission: a slot
        whose pair finished is refilled with the next request while the others keep decoding.
        Returns speech-token ids per request in vLLM id space (speech id + 1000), stop token
        included, like vLLM's output.token_ids."""
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            out: list = [None] * len(reqs)
            for i, toks, _ in self._run(reqs, poll_every):
                out[i] = toks
        torch.cuda.current_stream().wait_stream(self.stream)
        return out

    def generate_stream(self, reqs: list[Request], every: int = 32):
        """Streaming emission for the server caller (reference: openai_tts_server.py:392-528 consumes T3 tokens as
        they appear): a generator of (request index, new token ids in vLLM id space, finished) — each request's
        tokens since its last event, handed over every `every` decode steps from the device token ring.  The
        concatenation of a request's events is exactly what generate() returns for it."""
        self.stream.wait_stream(torch.cuda.current_stream())
        it = self._run(reqs, every, stream=True)
        while True:
            with torch.cuda.stream(self.stream):
                try:
                    ev = next(it)
                except StopIteration:
                    break
            yield ev
        torch.cuda.current_stream().wait_stream(self.stream)

    def generate_device(self, reqs: list[Request], tail_crop: int = 2, guard_caps: list[int] | None = None,
                        poll_every: int = 32) -> tuple[torch.Tensor, torch.Tensor]:
        """generate() with the vocoder hand-off done on the device (tts.py:448-486): returns
        (int64 [n, max_tokens] S3Gen-ready speech ids, int32 [n] counts), both CUDA tensors — no token
        leaves the GPU; the host only polls the `finished` flags."""
        n = len(reqs)
        width = max([r.params.max_tokens for r in reqs], default=1)
        dst = torch.zeros(max(n, 1), width, dtype=torch.int64, device=self.device)
        dst_len = torch.zeros(max(n, 1), dtype=torch.int32, device=self.device)
        caps = None if guard_caps is None else torch.tensor(guard_caps, dtype=torch.int32, device=self.device)
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for _ in self._run(reqs, poll_every, handoff=(dst, dst_len, caps, tail_crop)):
                pass
        torch.cuda.current_stream().wait_stream(self.stream)
        return dst[:n], dst_len[:n]

    def _generate(self, reqs: list[Request], poll_every: int, handoff=None) -> list[list[int]]:
        out: list = [None] * len(reqs)
        for i, toks, _ in self._run(reqs, poll_every, handoff=handoff):
            out[i] = toks
        return out

    def _run(self, reqs: list[Request], poll_every: int, handoff=None, stream: bool = False):
        """The scheduler loop (a generator).  KV blocks follow the sequences as in vLLM: a pair is admitted when the
        pool holds the blocks for its prefix plus one poll window (on top of what the running pairs need for their
        next window), every window tops each running row up to cdiv(tokens, 16) blocks, and when the pool runs dry
        the most recently admitted pair is preempted — its blocks go back, it rejoins the queue's head and, once
        re-admitted, its KV is recomputed over prefix + the tokens it had already emitted, after which it carries on
        from there (vLLM v1's preemption = recompute; nothing already handed out is generated again)."""
        n = len(reqs)
        if n == 0:
            return
        for r in reqs:
            if r.uid == 0:
                self._next_uid += 1
                r.uid = self._next_uid
        B = min(self.max_pairs, n)
        self._reset_slots(B)
        slot_req: list = [None] * B                            # request index living in each slot
        slot_held: dict = {}
        slot_len = [0] * B                                     # host upper bound of each slot's context length
        slot_seq = [0] * B                                     # admission order (the newest is preempted first)
        slot_emitted = [0] * B                                 # tokens of the slot's request already handed out (stream)
        history: dict = {}                                     # request index -> speech ids emitted before its preemption
        pending = list(range(n))
        admissions = 0

        def next_window_blocks() -> int:                       # blocks the running pairs take at their next top-up
            tot = 0
            for b in range(B):
                if slot_req[b] is not None:
                    cap = slot_held[b][2] + reqs[slot_req[b]].params.max_tokens
                    tot += 2 * max(0, -(-min(cap, slot_len[b] + poll_every) // BLOCK_SIZE) - len(self._rows[b]))
            return tot
        try:
            while True:
                free = [b for b in range(B) if slot_req[b] is None]
                new_slots, new_reqs = [], []
                while pending and free:
                    i = pending[0]
                    r = reqs[i]
                    if i in history:                           # recompute: same request, with what it had emitted
                        r = Request(r.text_ids, r.cond_emb, r.params, r.uid, r.penalize_prompt_ids, history[i])
                    p0 = COND_LEN + len(r.text_ids) + 1
                    first = min(p0 + r.params.max_tokens, p0 + len(r.history) + 1 + poll_every)
                    running = sum(q is not None for q in slot_req)
                    if 2 * -(-first // BLOCK_SIZE) + next_window_blocks() > self.pool.num_free():
                        if running == 0:
                            raise ValueError("KV pool too small for a single request")
                        break
                    b = free.pop(0)
                    slot_req[b] = pending.pop(0)
                    new_slots.append(b)
                    new_reqs.append(r)
                    admissions += 1
                    slot_seq[b] = admissions
                    # reserve now so the next fit check sees the pool shrink
                    P1, held1 = self._admit([r], [b], lookahead=poll_every)
                    slot_held[b] = (P1[0], held1, p0)          # rows prefilled, block lists, prefix length
                    slot_len[b] = P1[0] + 1
                    slot_emitted[b] = len(r.history)
                if new_slots:
                    P = [slot_held[b][0] for b in new_slots]
                    self._init_sampler_state(new_reqs, P, new_slots)
                    self.prefill(new_reqs, P, new_slots)
                active = [b for b in range(B) if slot_req[b] is not None]
                if not active:
                    break
                cap = lambda b: slot_held[b][2] + reqs[slot_req[b]].params.max_tokens      # noqa: E731
                remaining = max(cap(b) - slot_len[b] for b in active)
                if remaining > 0:
                    steps = min(poll_every, remaining)
                    while True:                                # blocks for this window: cdiv(tokens, 16) per row, slot order
                        want = []
                        for b in active:
                            tok = min(cap(b), slot_len[b] + steps)
                            want += [(b, tok), (B + b, tok)]
                        if self._grow(want):
                            break
                        if len(active) == 1:
                            raise ValueError("KV pool too small for a single request")
                        victim = max(active, key=lambda b: slot_seq[b])
                        n_v = int(self.sb.n_out[victim].item())  # what it has emitted so far stays emitted
                        history[slot_req[victim]] = self.sb.out_tokens[victim, :n_v].cpu().tolist()
                        self._park_slot(victim)
                        self._release(slot_held.pop(victim)[1])
                        pending.insert(0, slot_req[victim])
                        slot_req[victim] = None
                        active.remove(victim)
                        self.preemptions += 1
                    self.decode(B, steps, max(slot_len[b] for b in active) + steps)
                    for b in active:
                        slot_len[b] += steps
                fin = self.sb.finished[:B].cpu().tolist()      # the only host sync: every poll_every steps
                done = [b for b in active if fin[b]]
                if stream:
                    n_out = self.sb.n_out[:B].cpu().tolist()
                    for b in active:
                        lo = slot_emitted[b]
                        if n_out[b] > lo or fin[b]:
                            toks = (self.sb.out_tokens[b, lo:n_out[b]].cpu() + SPEECH_TOKEN_OFFSET).tolist()
                            slot_emitted[b] = n_out[b]
                            yield slot_req[b], toks, bool(fin[b])
                elif done and handoff is not None:          # retire on the device: post-process straight into the result rows
                    dst, dst_len, caps, crop = handoff
                    rows = torch.tensor([slot_req[b] for b in done], dtype=torch.int32, device=self.device)
                    ops.postprocess_tokens(self.sb, torch.tensor(done, dtype=torch.int32, device=self.device), dst, rows,
                                           dst_len, crop, None if caps is None else caps[rows.long()].contiguous(),
                                           eos=self.sb.stop_token)
                elif done:
                    n_out = self.sb.n_out[:B].cpu().tolist()
                    toks = self.sb.out_tokens[:B].cpu().numpy()
                    for b in done:
                        yield slot_req[b], (toks[b, :n_out[b]] + SPEECH_TOKEN_OFFSET).tolist(), True
                for b in done:
                    self._release(slot_held.pop(b)[1])
                    slot_req[b] = None
        finally:
            for _, held in slot_held.values():
                self._release(held)

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize()
            load().cbx_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
