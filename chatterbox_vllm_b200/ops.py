"""Per-kernel ops over torch CUDA tensors → libcbx_t3.so (C-ABI, include/cbx_t3.h).

torch is plumbing only (device memory + the current stream); all arithmetic happens in the
hand-written sm_100a kernels.  Every op validates device/dtype/contiguity and raises — there is
no eager/CPU fallback."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import (BLOCK_SIZE, HIDDEN, INTER, LOGITS_LD, PAGE_BYTES, SEEN_WORDS, SPEECH_VOCAB,
                   SamplerState, check, load)

BF16 = torch.bfloat16


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: torch.Tensor | None, dtype=None) -> int | None:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libcbx_t3 ops need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("libcbx_t3 ops need contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected dtype {dtype}, got {t.dtype}")
    return t.data_ptr()


def slot_mapping(block_table: torch.Tensor, req: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    out = torch.empty(req.numel(), dtype=torch.int64, device=req.device)
    check(load().cbx_slot_mapping(_p(block_table, torch.int32), block_table.stride(0), _p(req, torch.int32),
                                  _p(pos, torch.int32), _p(out), req.numel(), _stream()), "slot_mapping")
    return out


def rmsnorm(x: torch.Tensor, w: torch.Tensor, residual: torch.Tensor | None = None, eps: float = 1e-5):
    """out = norm(x)·w, or (residual += x in place; out = norm(residual)·w)."""
    assert x.shape[-1] == HIDDEN
    out = torch.empty_like(x)
    check(load().cbx_rmsnorm(_p(x, BF16), _p(residual, BF16), _p(w, BF16), _p(out), x.shape[0], eps, _stream()),
          "rmsnorm")
    return out


def pack_gate_up(wgu: torch.Tensor) -> torch.Tensor:
    out = torch.empty_like(wgu)
    check(load().cbx_pack_gate_up(_p(wgu, BF16), _p(out), wgu.shape[0] // 2, wgu.shape[1], _stream()), "pack_gate_up")
    return out


class Fp8Weight:
    """An nn.Linear weight [N, K] as FP8-e4m3 bytes + one fp32 scale per output feature (the optional
    weight-only FP8 path).  `scale_buf` is padded with 1.0 to a multiple of 128 entries, as the
    tcgen05 epilogue reads scales 32 at a time.  Every projection op below accepts one in place of a
    bf16 weight."""

    def __init__(self, q: torch.Tensor, scale_buf: torch.Tensor):
        self.q, self.scale_buf = q, scale_buf
        self.shape = q.shape

    @property
    def scale(self) -> torch.Tensor:
        return self.scale_buf[:self.q.shape[0]]

    def dequant(self) -> torch.Tensor:
        """fp32 [N, K] = q · scale — what the kernels multiply by (exactly, up to summation order)."""
        return self.q.view(torch.float8_e4m3fn).float() * self.scale[:, None]


def quantize_fp8_rows(W: torch.Tensor) -> Fp8Weight:
    """scale[n] = max|W[n,:]|/448, q = e4m3(W/scale) — on the GPU (cbx_quantize_fp8_rows)."""
    N, K = W.shape
    q = torch.empty(N, K, dtype=torch.uint8, device=W.device)
    sb = torch.ones((N + 127) // 128 * 128, dtype=torch.float32, device=W.device)
    check(load().cbx_quantize_fp8_rows(_p(W, BF16), _p(q), _p(sb), N, K, _stream()), "quantize_fp8_rows")
    return Fp8Weight(q, sb)


def _w8(W):
    return isinstance(W, Fp8Weight)


def linear(x: torch.Tensor, W, epilogue: int = 0, path: int = 0, out: torch.Tensor | None = None,
           ld: int | None = None):
    """y = x·Wᵀ ; epilogue=1: SiLU(gate)·up with W packed by pack_gate_up (y is [T, N/2]).
    ld: row stride of a freshly allocated result (elements) when N itself is not vector-aligned,
    e.g. the 8194-wide speech head; the returned tensor is the [T, N] view of it."""
    T, K = x.shape
    N = W.shape[0]
    assert W.shape[1] == K

    def run(dst, ldy, epi):
        if _w8(W):
            check(load().cbx_linear_fp8w(_p(x, BF16), _p(W.q, torch.uint8), _p(W.scale_buf, torch.float32), _p(dst, BF16),
                                         T, N, K, ldy, epi, path, _stream()), "linear_fp8w")
        else:
            check(load().cbx_linear(_p(x, BF16), _p(W, BF16), _p(dst, BF16), T, N, K, ldy, epi, path, _stream()), "linear")

    if ld is not None:
        assert out is None and epilogue == 0 and ld >= N
        buf = torch.empty(T, ld, dtype=BF16, device=x.device)
        run(buf, ld, 0)
        return buf[:, :N]
    if out is None:
        out = torch.empty(T, N // 2 if epilogue else N, dtype=BF16, device=x.device)
    run(out, out.stride(0), epilogue)
    return out


def rope_kv_append(qkv, pos, slot, cos_sin, kv_layer):
    check(load().cbx_rope_kv_append(_p(qkv, BF16), _p(pos, torch.int32), _p(slot, torch.int64), _p(cos_sin, BF16),
                                    _p(kv_layer), qkv.shape[0], _stream()), "rope_kv_append")


_attn_ws: dict = {}
_attn_planned: dict = {}     # (device, stream) -> T of the plan kept in that workspace


def paged_attn_decode(qkv, kv_layer, block_table, row_seq, kv_len, max_kv_len: int, out=None, plan: str = "build"):
    """plan: "build" (the call plans its page ranges itself), "make" (build the plan and keep it),
    "reuse" (a previous "make" call on this stream saw the same kv_len: skip the plan launch)."""
    T = qkv.shape[0]
    if out is None:
        out = torch.empty(T, HIDDEN, dtype=BF16, device=qkv.device)
    nbytes = load().cbx_attn_scratch_bytes(T, max_kv_len)
    key = (qkv.device, torch.cuda.current_stream().cuda_stream)
    ws = _attn_ws.get(key)
    if ws is None or ws[0].numel() < nbytes or ws[1].numel() < T:     # cached: tickets stay zero between calls
        ws = (torch.empty(max(nbytes, 16), dtype=torch.uint8, device=qkv.device),
              torch.zeros(max(T, 1024), dtype=torch.int32, device=qkv.device))
        _attn_ws[key] = ws
        _attn_planned.pop(key, None)
    fn = load().cbx_paged_attn_decode
    if plan == "make":
        rc = load().cbx_attn_plan_paged(_p(kv_len, torch.int32), T, _p(ws[0]), _stream(), _p(block_table, torch.int32),
                                        block_table.stride(0), _p(row_seq, torch.int32))
        if rc < 0:
            check(rc, "attn_plan")
        _attn_planned[key] = T if rc == 0 else 0
    if plan != "build" and _attn_planned.get(key) == T:
        fn = load().cbx_paged_attn_decode_planned
    check(fn(_p(qkv, BF16), _p(kv_layer), _p(block_table, torch.int32), block_table.stride(0), _p(row_seq, torch.int32),
             _p(kv_len, torch.int32), _p(out), _p(ws[0]), _p(ws[1], torch.int32), T, max_kv_len, _stream()),
          "paged_attn_decode")
    return out


def paged_attn_decode_static(qkv, kv_layer, block_table, row_seq, kv_len, max_kv_len: int, out):
    """cbx_paged_attn_decode_static: the allocation-free form (scratch lives inside the library; 5 <= T <= 4095).
    The caller passes `out`; nothing is allocated here either."""
    check(load().cbx_paged_attn_decode_static(_p(qkv, BF16), _p(kv_layer), _p(block_table, torch.int32),
                                              block_table.stride(0), _p(row_seq, torch.int32), _p(kv_len, torch.int32),
                                              _p(out, BF16), qkv.shape[0], max_kv_len, _stream()),
          "paged_attn_decode_static")
    return out


def prefill_items(row_starts: list[int], lens: list[int], bt_rows: list[int]) -> torch.Tensor:
    """Host helper: int32 [n,4] items {row0, n_q, block-table row, query block} for sequences whose
    prefix rows start at row_starts[i] (positions 0..lens[i]-1), longest items first (they are dealt
    round-robin to the persistent CTAs)."""
    it = []
    for r0, n, br in zip(row_starts, lens, bt_rows):
        for qb in range(-(-n // BLOCK_SIZE)):
            it.append((r0 + qb * BLOCK_SIZE, min(BLOCK_SIZE, n - qb * BLOCK_SIZE), br, qb))
    it.sort(key=lambda t: -t[3])
    return torch.tensor(it, dtype=torch.int32).view(-1, 4)


def paged_attn_prefill(qkv, kv_layer, block_table, items, out=None):
    T = qkv.shape[0]
    if out is None:
        out = torch.zeros(T, HIDDEN, dtype=BF16, device=qkv.device)
    check(load().cbx_paged_attn_prefill(_p(qkv, BF16), _p(kv_layer), _p(block_table, torch.int32), block_table.stride(0),
                                        _p(items, torch.int32), items.shape[0], _p(out), _stream()), "paged_attn_prefill")
    return out


def layer_qkv(res, ln_w, wqkv, cos_sin, pos, slot, kv_layer, eps: float = 1e-5, qkv=None, h=None):
    """RMSNorm → QKV → RoPE → KV append.  Returns the qkv buffer [T,3072] whose first 1024 columns hold
    the rotated q (the k/v columns are not written: they went to the cache)."""
    T = res.shape[0]
    if qkv is None:
        qkv = torch.zeros(T, 3 * HIDDEN, dtype=BF16, device=res.device)
    if h is None and T > 4:
        h = torch.empty(T, HIDDEN, dtype=BF16, device=res.device)
    if _w8(wqkv):
        check(load().cbx_layer_qkv_fp8w(_p(res, BF16), _p(ln_w, BF16), _p(wqkv.q, torch.uint8),
                                        _p(wqkv.scale_buf, torch.float32), _p(cos_sin, BF16), _p(pos, torch.int32),
                                        _p(slot, torch.int64), _p(qkv), _p(kv_layer), _p(h), T, eps, _stream()),
              "layer_qkv_fp8w")
        return qkv
    check(load().cbx_layer_qkv(_p(res, BF16), _p(ln_w, BF16), _p(wqkv, BF16), _p(cos_sin, BF16), _p(pos, torch.int32),
                               _p(slot, torch.int64), _p(qkv), _p(kv_layer), _p(h), T, eps, _stream()), "layer_qkv")
    return qkv


_block_ws: dict = {}
_mlp_ws: dict = {}


def layer_attn_block(res, ln_w, wqkv, wo, cos_sin, pos, slot, kv_layer, block_table, row_seq, kv_len, max_kv_len: int,
                     eps: float = 1e-5, q_out=None):
    """cbx_layer_attn_block: norm + QKV + RoPE + KV append + paged attention + O-proj + residual in one launch
    (T <= 4 rows).  Updates `res` and the KV pages in place."""
    T = res.shape[0]
    ws = _block_ws.get(res.device)
    if ws is None:
        ws = torch.zeros(load().cbx_layer_attn_block_scratch_bytes(), dtype=torch.uint8, device=res.device)
        _block_ws[res.device] = ws
    check(load().cbx_layer_attn_block(_p(res, BF16), _p(ln_w, BF16), _p(wqkv, BF16), _p(wo, BF16), _p(cos_sin, BF16),
                                      _p(pos, torch.int32), _p(slot, torch.int64), _p(kv_layer), _p(block_table, torch.int32),
                                      block_table.stride(0), _p(row_seq, torch.int32), _p(kv_len, torch.int32), _p(ws),
                                      _p(q_out, BF16), T, max_kv_len, eps, _stream()), "layer_attn_block")
    return res


def layer_mlp_block(res, ln_w, wgu_packed, wd_t, eps: float = 1e-5):
    """cbx_layer_mlp_block: norm + gate_up + SiLU·mul + down-proj + residual in one launch (T <= 4 rows, bf16 weights).
    wd_t = Wd.t().contiguous() ([4096, 1024]).  Updates `res` in place."""
    T = res.shape[0]
    assert tuple(wd_t.shape) == (4096, HIDDEN) and tuple(wgu_packed.shape) == (8192, HIDDEN)
    ws = _mlp_ws.get(res.device)
    if ws is None:
        ws = torch.zeros(load().cbx_layer_mlp_block_scratch_bytes(), dtype=torch.uint8, device=res.device)
        _mlp_ws[res.device] = ws
    check(load().cbx_layer_mlp_block(_p(res, BF16), _p(ln_w, BF16), _p(wgu_packed, BF16), _p(wd_t, BF16), _p(ws), T, eps,
                                     _stream()), "layer_mlp_block")
    return res


def linear_residual(x, W, res):
    """res += x·Wᵀ in place (N = 1024)."""
    T, K = x.shape
    assert tuple(W.shape) == (HIDDEN, K) and res.shape == (T, HIDDEN)
    if _w8(W):
        check(load().cbx_linear_residual_fp8w(_p(x, BF16), _p(W.q, torch.uint8), _p(W.scale_buf, torch.float32),
                                              _p(res, BF16), T, K, _stream()), "linear_residual_fp8w")
        return res
    check(load().cbx_linear_residual(_p(x, BF16), _p(W, BF16), _p(res, BF16), T, K, _stream()), "linear_residual")
    return res


def linear_partials(x, W, splits: int):
    """fp32 [splits, T, 1024] split-K partial products of x·Wᵀ."""
    T, K = x.shape
    part = torch.empty(splits, T, HIDDEN, dtype=torch.float32, device=x.device)
    if _w8(W):
        check(load().cbx_linear_partials_fp8w(_p(x, BF16), _p(W.q, torch.uint8), _p(W.scale_buf, torch.float32), _p(part),
                                              T, K, splits, _stream()), "linear_partials_fp8w")
        return part
    check(load().cbx_linear_partials(_p(x, BF16), _p(W, BF16), _p(part), T, K, splits, _stream()), "linear_partials")
    return part


def rmsnorm_reduce(part, res, w, eps: float = 1e-5):
    """res = bf16(bf16(Σ part) + res) in place; returns norm(res)·w."""
    splits, T, _ = part.shape
    out = torch.empty_like(res)
    check(load().cbx_rmsnorm_reduce(_p(part, torch.float32), splits, _p(res, BF16), _p(w, BF16), _p(out), T, eps,
                                    _stream()), "rmsnorm_reduce")
    return out


def layer_gate_up(res, ln_w, wgu_packed, eps: float = 1e-5, act=None, h=None):
    T = res.shape[0]
    if act is None:
        act = torch.empty(T, INTER, dtype=BF16, device=res.device)
    if h is None and T > 4:
        h = torch.empty(T, HIDDEN, dtype=BF16, device=res.device)
    if _w8(wgu_packed):
        check(load().cbx_layer_gate_up_fp8w(_p(res, BF16), _p(ln_w, BF16), _p(wgu_packed.q, torch.uint8),
                                            _p(wgu_packed.scale_buf, torch.float32), _p(act), _p(h), T, eps, _stream()),
              "layer_gate_up_fp8w")
        return act
    check(load().cbx_layer_gate_up(_p(res, BF16), _p(ln_w, BF16), _p(wgu_packed, BF16), _p(act), _p(h), T, eps,
                                   _stream()), "layer_gate_up")
    return act


def head_logits(res, norm_w, whead, eps: float = 1e-5, logits=None, h=None):
    T = res.shape[0]
    if logits is None:
        logits = torch.zeros(T, LOGITS_LD, dtype=BF16, device=res.device)
    if h is None and T > 4:
        h = torch.empty(T, HIDDEN, dtype=BF16, device=res.device)
    if _w8(whead):
        check(load().cbx_head_logits_fp8w(_p(res, BF16), _p(norm_w, BF16), _p(whead.q, torch.uint8),
                                          _p(whead.scale_buf, torch.float32), _p(logits), _p(h), T, eps, _stream()),
              "head_logits_fp8w")
        return logits
    check(load().cbx_head_logits(_p(res, BF16), _p(norm_w, BF16), _p(whead, BF16), _p(logits), _p(h), T, eps,
                                 _stream()), "head_logits")
    return logits


class SamplerBuffers:
    """Device-resident per-pair sampler state (cbx_sampler_state)."""

    def __init__(self, B: int, out_stride: int, device, *, stop_token: int = 6562, cfg_scale: float = 0.5):
        z = lambda *s, dt=torch.int32: torch.zeros(*s, dtype=dt, device=device)  # noqa: E731
        self.B, self.out_stride = B, out_stride
        self.last_tok, self.out_tokens, self.n_out = z(B), z(B, out_stride), z(B)
        self.max_out, self.finished, self.seq_len = z(B), z(B), z(2 * B)
        self.seen = z(B, SEEN_WORDS)
        self.temperature, self.top_p = z(B, dt=torch.float32), torch.ones(B, dtype=torch.float32, device=device)
        self.min_p, self.rep_pen = z(B, dt=torch.float32), torch.ones(B, dtype=torch.float32, device=device)
        self.top_k, self.ignore_eos = z(B), z(B)
        self.seed, self.uid = z(B, dt=torch.int64), z(B, dt=torch.int64)
        self.n_finished = z(1)
        self.stop_token, self.cfg_scale = stop_token, cfg_scale

    def struct(self) -> SamplerState:
        s = SamplerState()
        for name in ("last_tok", "out_tokens", "n_out", "max_out", "finished", "seq_len", "seen", "temperature",
                     "top_p", "min_p", "rep_pen", "top_k", "ignore_eos", "seed", "uid", "n_finished"):
            setattr(s, name, getattr(self, name).data_ptr())
        s.out_stride, s.stop_token, s.cfg_scale = self.out_stride, self.stop_token, self.cfg_scale
        return s


def cfg_sample(logits: torch.Tensor, sb: SamplerBuffers, B: int) -> None:
    assert logits.shape[1] == LOGITS_LD and logits.shape[0] >= 2 * B
    st = sb.struct()
    check(load().cbx_cfg_sample(_p(logits, BF16), C.byref(st), B, _stream()), "cfg_sample")


def decode_prepare(speech_emb, sb: SamplerBuffers, block_table, B: int):
    dev = speech_emb.device
    x = torch.empty(2 * B, HIDDEN, dtype=BF16, device=dev)
    pos = torch.empty(2 * B, dtype=torch.int32, device=dev)
    slot = torch.empty(2 * B, dtype=torch.int64, device=dev)
    kv_len = torch.empty(2 * B, dtype=torch.int32, device=dev)
    check(load().cbx_decode_prepare(_p(speech_emb, BF16), _p(sb.last_tok), _p(sb.finished), _p(sb.seq_len),
                                    _p(block_table, torch.int32), block_table.stride(0), _p(x), _p(pos), _p(slot),
                                    _p(kv_len), B, _stream()), "decode_prepare")
    return x, pos, slot, kv_len


def prefill_embed(desc, cond_emb, text_emb, text_pos, speech_emb, speech_pos, out=None):
    T = desc.shape[0]
    if out is None:
        out = torch.empty(T, HIDDEN, dtype=BF16, device=desc.device)
    check(load().cbx_prefill_embed(_p(desc, torch.int32), _p(cond_emb, torch.float32), _p(text_emb, BF16),
                                   _p(text_pos, BF16), _p(speech_emb, BF16), _p(speech_pos, BF16), _p(out), T,
                                   _stream()), "prefill_embed")
    return out


def prefill_rows(seq, text_ids, hist, T: int, desc=None, pos=None, row_seq=None, kv_len=None):
    """cbx_prefill_rows: per-row descriptors / positions / block-table rows / context lengths of a prefill step from one
    int32 [n_seq, 8] record per sequence (see include/cbx_t3.h).  Returns (desc [T,4], pos, row_seq, kv_len)."""
    dev = seq.device
    new = lambda *shape: torch.empty(*shape, dtype=torch.int32, device=dev)  # noqa: E731
    desc = new(T, 4) if desc is None else desc
    pos = new(T) if pos is None else pos
    row_seq = new(T) if row_seq is None else row_seq
    kv_len = new(T) if kv_len is None else kv_len
    check(load().cbx_prefill_rows(_p(seq, torch.int32), seq.shape[0], _p(text_ids, torch.int32), _p(hist, torch.int32), T,
                                  _p(desc), _p(pos), _p(row_seq), _p(kv_len), _stream()), "prefill_rows")
    return desc, pos, row_seq, kv_len


def postprocess_tokens(sb: "SamplerBuffers", slots: torch.Tensor, dst: torch.Tensor, dst_rows: torch.Tensor,
                       dst_len: torch.Tensor, tail_crop: int = 2, guard_cap: torch.Tensor | None = None,
                       sos: int = 6561, eos: int = 6562) -> None:
    """Device-side T3 -> vocoder hand-off for the sampler slots `slots` (int32 [n]): stop cut, tail crop,
    SOS/EOS framing, < 6561 filter, guard cap; int64 ids into dst[dst_rows[i]], counts into dst_len."""
    n = slots.numel()
    check(load().cbx_postprocess_tokens(_p(sb.out_tokens, torch.int32), _p(sb.n_out, torch.int32), sb.out_stride,
                                        _p(slots, torch.int32), n, sos, eos, tail_crop, _p(guard_cap, torch.int32),
                                        _p(dst, torch.int64), dst.stride(0), _p(dst_rows, torch.int32),
                                        _p(dst_len, torch.int32), _stream()), "postprocess_tokens")


class BlockPool:
    """Host block pool (C++ in libcbx_t3.so) with vLLM's allocation policy."""

    def __init__(self, num_blocks: int):
        self._h = load().cbx_pool_create(num_blocks)
        if not self._h:
            raise RuntimeError(load().cbx_last_error().decode())

    def num_free(self) -> int:
        return load().cbx_pool_num_free(self._h)

    def alloc(self, n: int) -> list[int]:
        buf = (C.c_int32 * max(n, 1))()
        rc = load().cbx_pool_alloc(self._h, n, buf)
        if rc != 0:
            raise ValueError(load().cbx_last_error().decode())
        return list(buf[:n])

    def free(self, ids: list[int]) -> None:
        buf = (C.c_int32 * max(len(ids), 1))(*ids)
        check(load().cbx_pool_free(self._h, buf, len(ids)), "pool_free")

    def __del__(self):
        try:
            if self._h:
                load().cbx_pool_destroy(self._h)
                self._h = None
        except Exception:
            pass


__all__ = ["slot_mapping", "rmsnorm", "pack_gate_up", "linear", "rope_kv_append", "paged_attn_decode",
           "layer_qkv", "prefill_items", "paged_attn_prefill", "linear_residual", "linear_partials", "rmsnorm_reduce", "layer_gate_up", "head_logits",
           "SamplerBuffers", "cfg_sample", "decode_prepare", "prefill_embed", "BlockPool", "BLOCK_SIZE", "HIDDEN",
           "INTER", "LOGITS_LD", "PAGE_BYTES", "SEEN_WORDS", "SPEECH_VOCAB", "_lib"]
