"""GPU parity tests of the optional FP8-e4m3 weight path (north star: "an optional FP8-e4m3
per-channel weight path is reported separately with its own tolerance"; SURVEY §8(d) cfg 5).

Three kinds of evidence, all through the C-ABI:
  * the quantiser is BIT-EXACT against torch's float8_e4m3fn cast (integer/byte work);
  * on weights whose scales are powers of two the dequantised weights are exact bf16 numbers, so the
    FP8 kernels must reproduce the (already oracle-checked) bf16 kernels: bit for bit on the tcgen05
    path — same operand bits in shared memory, same MMA sequence, exact scale multiply — which pins the
    on-chip e4m3 -> bf16 conversion and its SWIZZLE_128B placement; within summation-order noise on
    the GEMV path;
  * on ordinary weights the kernels and the whole engine are held to the oracle run on the
    dequantised weights (q * scale in fp32), same bars as the bf16 path; the distance between the
    FP8 model and the bf16 model is REPORTED (and bounded by the tolerance SURVEY states)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import sampler_ref
from oracle.t3_ref import (T3ModelConfig, T3Oracle, T3Weights, apply_rope, fp8_e4m3_rows, rmsnorm, rope_cos_sin,
                           with_fp8_projections)
from test_gpu_engine import _requests, assert_stream_parity
from test_gpu_kernels import ULP, _unswizzled_k, assert_close_bf16

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200 import ops as _ops
    _ops._lib.load()
    return _ops


def dev(t):
    return t.cuda().contiguous()


def _pow2_weight(N, K, seed):
    """bf16 [N, K] whose per-row e4m3 quantisation is exact: random finite e4m3 bytes times 2^k, one
    |448| per row so that scale = 2^k exactly."""
    g = torch.Generator().manual_seed(seed)
    b = torch.randint(0, 256, (N, K), generator=g, dtype=torch.int32)
    b = torch.where((b & 0x7f) == 0x7f, b ^ 0x08, b).to(torch.uint8)           # no NaN encodings
    b[:, 0] = 0x7e                                                            # +448
    k = torch.randint(-13, -8, (N,), generator=g)
    scale = torch.pow(2.0, k.float())
    W = (b.view(torch.float8_e4m3fn).float() * scale[:, None]).to(BF16)
    assert torch.equal(W.float(), b.view(torch.float8_e4m3fn).float() * scale[:, None])
    return W, b, scale


def _ref(x, Wdq):
    return (x.double() @ Wdq.double().T).float()


# ------------------------------------------------------------------ quantiser: bit-exact
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 4096), (8194, 1024), (7, 64)])
def test_quantize_fp8_rows_bit_exact(ops, N, K):
    g = torch.Generator().manual_seed(N)
    W = (torch.randn(N, K, generator=g) * 0.02).to(BF16)
    W[min(3, N - 1)] = 0                                                      # a zero row keeps scale 1
    q8 = ops.quantize_fp8_rows(dev(W))
    q, scale = fp8_e4m3_rows(W)
    assert torch.equal(q8.scale.cpu(), scale), "scales must be bit-identical"
    assert torch.equal(q8.q.cpu(), q.view(torch.uint8)), "e4m3 bytes must be bit-identical"
    assert q8.scale_buf.numel() % 128 == 0 and (q8.scale_buf[N:] == 1).all()
    assert torch.equal(q8.dequant().cpu(), q.float() * scale[:, None])
    # exactly representable weights survive the round trip
    Wp, b, sc = _pow2_weight(64, K, 5)
    qp = ops.quantize_fp8_rows(dev(Wp))
    assert torch.equal(qp.q.cpu(), b) and torch.equal(qp.scale.cpu(), sc)


# ------------------------------------------------------------------ GEMV regime (T <= 4)
@pytest.mark.parametrize("T", [1, 2, 3, 4])
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 4096), (8194, 1024)])
def test_linear_fp8w_gemv(ops, T, N, K):
    g = torch.Generator().manual_seed(N + T)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(N, K, generator=g) * 0.05).to(BF16)
    q8 = ops.quantize_fp8_rows(dev(W))
    y = ops.linear(dev(x), q8, path=1, ld=8256 if N == 8194 else None)
    assert_close_bf16(y, _ref(x, q8.dequant().cpu()), ulps=1.0, atol=2e-3, what=f"fp8w gemv T={T} N={N} K={K}")
    # quantisation itself moves the result far more than that: the test above is not vacuous
    assert (y.float().cpu() - _ref(x, W.float())).abs().max() > 8 * 2e-3


# ------------------------------------------------------------------ tcgen05 regime: converter pinned bit for bit
@pytest.mark.parametrize("T", [5, 64, 200, 256, 300, 2500])
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 1024), (1024, 4096), (8194, 1024)])
def test_linear_fp8w_gemm_bit_exact_on_pow2_scales(ops, T, N, K):
    g = torch.Generator().manual_seed(N + 7 * T)
    x = torch.randn(T, K, generator=g).to(BF16)
    W, b, sc = _pow2_weight(N, K, N + T)
    q8 = ops.quantize_fp8_rows(dev(W))
    assert torch.equal(q8.q.cpu(), b)
    ld = 8256 if N == 8194 else N
    out8 = torch.full((T, ld), 7.0, dtype=BF16).cuda()
    out16 = torch.full((T, ld), 7.0, dtype=BF16).cuda()
    ops.linear(dev(x), q8, path=2, out=out8)
    ops.linear(dev(x), dev(W), path=2, out=out16)
    assert_close_bf16(out8[:, :N], _ref(x, W.float()), ulps=1.0, atol=2e-3, what=f"fp8w gemm T={T} N={N} K={K}")
    assert torch.equal(out8, out16), "same operand bits + exact power-of-two scale => same output bits (and untouched padding)"


@pytest.mark.parametrize("T", [7, 128, 256, 1000])
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 4096), (8194, 1024)])
def test_linear_fp8w_gemm_random_scales(ops, T, N, K):
    g = torch.Generator().manual_seed(N + 11 * T)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(N, K, generator=g) * 0.05 * (0.25 + torch.rand(N, 1, generator=g))).to(BF16)
    q8 = ops.quantize_fp8_rows(dev(W))
    y = ops.linear(dev(x), q8, path=2, ld=8256 if N == 8194 else None)
    assert_close_bf16(y, _ref(x, q8.dequant().cpu()), ulps=1.0, atol=2e-3, what=f"fp8w gemm T={T} N={N} K={K}")


@pytest.mark.parametrize("T,path", [(2, 1), (4, 1), (6, 2), (256, 2), (700, 2)])
def test_linear_fp8w_silu_mul_epilogue(ops, T, path):
    g = torch.Generator().manual_seed(T)
    x = torch.randn(T, 1024, generator=g).to(BF16)
    Wgu = (torch.randn(8192, 1024, generator=g) * 0.05).to(BF16)
    q8 = ops.quantize_fp8_rows(ops.pack_gate_up(dev(Wgu)))            # scales travel in packed row order
    q, s = fp8_e4m3_rows(Wgu)
    y = ops.linear(dev(x), q8, epilogue=1, path=path)
    gu = _ref(x, q.float() * s[:, None]).to(BF16).float()
    want = torch.nn.functional.silu(gu[:, :4096]) * gu[:, 4096:]
    assert_close_bf16(y, want, ulps=3.0, atol=3e-3, what=f"fp8w silu_mul T={T}")


# ------------------------------------------------------------------ fused layer pieces
@pytest.mark.parametrize("T", [1, 2, 4, 6, 200, 1600])
def test_layer_qkv_fp8w_norm_rope_append(ops, T):
    cfg = T3ModelConfig()
    g = torch.Generator().manual_seed(100 + T)
    res = (torch.randn(T, 1024, generator=g) * 1.5).to(BF16)
    ln = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    W = (torch.randn(3072, 1024, generator=g) * 0.03).to(BF16)
    pos = torch.randint(0, 900, (T,), generator=g, dtype=torch.int32)
    n_blocks = T + 2
    blk = torch.randperm(n_blocks - 1, generator=g)[:T] + 1
    slot = (blk * 16 + torch.randint(0, 16, (T,), generator=g)).to(torch.int64)
    if T > 1:
        slot[T - 1] = -1
    kv = torch.zeros(n_blocks * 65536, dtype=torch.uint8).cuda()
    cs = rope_cos_sin(cfg, 1024)
    q8 = ops.quantize_fp8_rows(dev(W))
    qkv = ops.layer_qkv(dev(res), dev(ln), q8, dev(cs.to(BF16)), dev(pos), dev(slot), kv)
    h = rmsnorm(res.float(), ln, 1e-5, "bf16")
    y = _ref(h, q8.dequant().cpu()).to(BF16).float()
    want_q = apply_rope(y[:, :1024].view(T, 16, 64), pos.long(), cs, "bf16").reshape(T, 1024)
    want_k = apply_rope(y[:, 1024:2048].view(T, 16, 64), pos.long(), cs, "bf16").reshape(T, 1024)
    want_v = y[:, 2048:]
    tol = dict(ulps=3.0, atol=0.03)
    assert_close_bf16(qkv[:, :1024], want_q, what="fp8w fused q", **tol)
    pages = kv.cpu().view(-1, 65536)
    for t in range(min(T, 40)):
        if slot[t] < 0:
            continue
        b, tok = int(slot[t]) // 16, int(slot[t]) % 16
        page = pages[b].view(torch.bfloat16)
        for hd in range(16):
            assert_close_bf16(_unswizzled_k(page, hd, tok), want_k[t, hd * 64:hd * 64 + 64], what="fp8w fused K", **tol)
            vrow = page[16384 + (hd * 16 + tok) * 64:16384 + (hd * 16 + tok) * 64 + 64]
            assert_close_bf16(vrow, want_v[t, hd * 64:hd * 64 + 64], ulps=1.01, atol=0.02, what="fp8w fused V")
    if T > 1:
        assert not pages[0].any(), "the null block must stay untouched"


@pytest.mark.parametrize("T", [1, 3, 4, 5, 256, 2000])
@pytest.mark.parametrize("K", [1024, 4096])
def test_linear_residual_fp8w_in_place(ops, T, K):
    g = torch.Generator().manual_seed(T * 7 + K)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(1024, K, generator=g) * 0.03).to(BF16)
    res = (torch.randn(T, 1024, generator=g) * 2).to(BF16)
    q8 = ops.quantize_fp8_rows(dev(W))
    got = ops.linear_residual(dev(x), q8, dev(res.clone()))
    y = _ref(x, q8.dequant().cpu()).to(BF16).float()
    want = (y + res.float()).to(BF16).float()
    err = (got.float().cpu() - want).abs()
    tol = 2.0 * ULP * y.abs() + 1.01 * ULP * want.abs() + 1e-6
    assert (err <= tol).all(), f"max err {err.max():.3e}"


@pytest.mark.parametrize("T,K,splits", [(5, 1024, 4), (256, 1024, 4), (256, 4096, 4), (300, 4096, 2)])
def test_split_k_partials_fp8w(ops, T, K, splits):
    g = torch.Generator().manual_seed(T + K)
    x = torch.randn(T, K, generator=g).to(BF16)
    W, _, _ = _pow2_weight(1024, K, T + K)
    q8 = ops.quantize_fp8_rows(dev(W))
    p8 = ops.linear_partials(dev(x), q8, splits)
    p16 = ops.linear_partials(dev(x), dev(W), splits)
    assert torch.equal(p8, p16), "split-K partials: same bits as the bf16 kernel on power-of-two scales"
    Wr = (torch.randn(1024, K, generator=g) * 0.03).to(BF16)
    qr = ops.quantize_fp8_rows(dev(Wr))
    pr = ops.linear_partials(dev(x), qr, splits).sum(0).cpu()
    want = _ref(x, qr.dequant().cpu())
    assert (pr - want).abs().max() <= 2e-3 + 1e-3 * want.abs().max()


@pytest.mark.parametrize("T", [2, 4, 6, 256])
def test_head_and_gate_up_fp8w_fused_norm(ops, T):
    g = torch.Generator().manual_seed(T + 77)
    res = (torch.randn(T, 1024, generator=g) * 1.2).to(BF16)
    nw = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    Wh = (torch.randn(8194, 1024, generator=g) * 0.05).to(BF16)
    q8 = ops.quantize_fp8_rows(dev(Wh))
    lg = ops.head_logits(dev(res), dev(nw), q8)
    h = rmsnorm(res.float(), nw, 1e-5, "bf16")
    assert_close_bf16(lg[:, :8194], _ref(h, q8.dequant().cpu()), ulps=2.0, atol=0.02, what=f"fp8w head T={T}")
    Wgu = (torch.randn(8192, 1024, generator=g) * 0.03).to(BF16)
    act = ops.layer_gate_up(dev(res), dev(nw), ops.quantize_fp8_rows(ops.pack_gate_up(dev(Wgu))))
    q, s = fp8_e4m3_rows(Wgu)
    gu = _ref(h, q.float() * s[:, None]).to(BF16).float()
    want = torch.nn.functional.silu(gu[:, :4096]) * gu[:, 4096:]
    assert_close_bf16(act, want, ulps=3.0, atol=3e-3, what=f"fp8w fused gate_up T={T}")


# ------------------------------------------------------------------ the whole engine on FP8 weights
def _mk_engine8(n_layers, max_pairs, **kw):
    from chatterbox_vllm_b200.engine import T3Engine
    cfg = T3ModelConfig(n_layers=n_layers)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    eng = T3Engine(w.state_dict(), device="cuda:0", max_pairs=max_pairs, max_model_len=256, weight_dtype="fp8_e4m3", **kw)
    return cfg, w, with_fp8_projections(w), eng


@pytest.mark.parametrize("use_graph", [False, True])
def test_engine_fp8_greedy_stream_matches_fp8_oracle(ops, use_graph):
    """Same bar as the bf16 engine, against the oracle run on the dequantised weights: GEMV regime
    (1 pair), tcgen05 regime (3 pairs), graph replay and eager."""
    n_steps = 24
    cfg, w, w8, eng = _mk_engine8(2, 4, use_graph=use_graph)
    assert eng.weight_dtype == "fp8_e4m3" and eng.projection_weight_bytes() < 0.52 * 2 * (2 * 16_779_264 + 8194 * 1024)
    reqs = _requests(3, n_steps, temperature=0.0, repetition_penalty=2.0)
    got = eng.generate(reqs, poll_every=5)
    assert_stream_parity(got, w8, reqs, eng.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
    assert eng.generate(reqs) == got and eng.pool.num_free() == eng.num_blocks - 1
    one = eng.generate(reqs[:1])
    assert_stream_parity(one, w8, reqs[:1], eng.cfg_scale, n_steps, temperature=0.0, repetition_penalty=2.0)
    eng.close()


def test_engine_fp8_batch32_split_k_and_sampling(ops):
    n_steps = 10
    cfg, w, w8, eng = _mk_engine8(2, 32)
    reqs = _requests(32, n_steps, temperature=0.0, repetition_penalty=2.0)
    got = eng.generate(reqs, poll_every=4)
    assert eng.generate(reqs, poll_every=3) == got, "same batch must give the same bits"
    assert_stream_parity(got, w8, reqs, eng.cfg_scale, n_steps, first_exact=0, temperature=0.0, repetition_penalty=2.0)
    # sampled streams: deterministic under the same Philox counters, all tokens valid speech ids
    sreqs = _requests(8, 12, temperature=0.5, top_p=0.5, repetition_penalty=2.0)
    s1 = eng.generate(sreqs)
    assert s1 == eng.generate(sreqs) and all(1000 <= t < 1000 + 8194 for r in s1 for t in r)
    eng.close()


def test_engine_fp8_vs_bf16_model_distance_30layers(ops):
    """The FP8 path's OWN tolerance.  Per-channel e4m3 keeps 4 significant bits of every weight (relative
    rounding error uniform in +-2^-4, 3.6 % rms), so each projection output moves by ~3.6 % and 30 layers x 4
    projections compound: on the sharp random-init model (speech_head x 16, logit std 11) the oracle puts the
    FP8 model's head logits at rel-L2 0.17 / max-abs 7.4 from the bf16 model.  SURVEY 8(d)'s "max-abs <= 0.5"
    guess is therefore not attainable by ANY per-channel e4m3 weight path on Gaussian weights; the bar here
    is rel-L2 <= 0.25 against the bf16 model (reported with max-abs and greedy agreement), while the KERNELS
    are held to the FP8 oracle (dequantised weights) with the bf16 path's own 30-layer bar."""
    from chatterbox_vllm_b200.engine import T3Engine
    n_steps = 16
    cfg = T3ModelConfig(n_layers=30)
    w = T3Weights.random(cfg, seed=0, head_scale=16.0, norm_jitter=0.1)
    reqs = _requests(2, n_steps, temperature=0.0, repetition_penalty=2.0)
    e16 = T3Engine(w.state_dict(), device="cuda:0", max_pairs=2, max_model_len=256)
    t16 = e16.generate(reqs)
    e16.close()
    e8 = T3Engine(w.state_dict(), device="cuda:0", max_pairs=2, max_model_len=256, weight_dtype="fp8_e4m3")
    t8 = e8.generate(reqs)
    e8.close()
    w8 = with_fp8_projections(w)
    assert_stream_parity(t8, w8, reqs, 0.5, n_steps, first_exact=0, min_agree=0.8, temperature=0.0, repetition_penalty=2.0)
    # model distance on the oracle side (fp32 accumulation, bf16 storage): prefill logits of both models
    o16, o8 = T3Oracle(w, "bf16"), T3Oracle(w8, "bf16")
    worst_abs = worst_rel = 0.0
    for r in reqs:
        l16 = o16.prefill_pair(o16.new_pair(), r.cond_emb, r.text_ids, 0.5)
        l8 = o8.prefill_pair(o8.new_pair(), r.cond_emb, r.text_ids, 0.5)
        worst_abs = max(worst_abs, float((l16 - l8).abs().max()))
        worst_rel = max(worst_rel, float((l16 - l8).norm() / l16.norm()))
    agree = sum(a == b for x, y in zip(t8, t16) for a, b in zip(x, y)) / (len(reqs) * n_steps)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/fp8_model_distance.json", "w") as f:
        json.dump({"head_logit_rel_l2_fp8_vs_bf16_model": worst_rel, "head_logit_max_abs_fp8_vs_bf16_model": worst_abs,
                   "greedy_agreement_fp8_vs_bf16_engine": agree, "n_steps": n_steps,
                   "variant": "sharp (speech_head x16), 30 layers, random-init"}, f)
    assert worst_rel <= 0.25, f"FP8 model moved the head logits by rel-L2 {worst_rel:.3f} (> 0.25)"
