"""GPU parity tests, kernel by kernel, through the C-ABI (chatterbox_vllm_b200.ops) against
the CPU oracle.  Integer work is bit-exact; floating point has its tolerance written at the
assert.  bf16 has an 8-bit significand: 1 ulp = 2^-8 relative."""
import numpy as np
import random

import pytest
import torch

from oracle import paging_ref, sampler_ref
from oracle.t3_ref import T3ModelConfig, T3Oracle, T3Weights, apply_rope, rmsnorm, rope_cos_sin

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
ULP = 2.0 ** -7


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200 import ops as _ops
    _ops._lib.load()
    return _ops


def dev(t):
    return t.cuda()


def assert_close_bf16(got, want, ulps=2.0, atol=1e-6, what=""):
    got, want = got.float().cpu(), want.float().cpu()
    err = (got - want).abs()
    tol = ulps * ULP * want.abs() + atol
    bad = err > tol
    assert not bad.any(), f"{what}: {int(bad.sum())} elements off; max err {err.max():.3e} (|want| max {want.abs().max():.3e})"


# ------------------------------------------------------------------ integer work: bit-exact
def test_slot_mapping_bit_exact(ops):
    rng = np.random.default_rng(0)
    for occupancy in (0.25, 0.5, 0.95):
        pool = paging_ref.BlockPool(4096)
        bt = paging_ref.BlockTable(pool, max_seqs=64, max_model_len=1000)
        lens = rng.integers(1, int(1000 * occupancy) + 2, size=64)
        for r, n in enumerate(lens):
            bt.ensure(r, int(n))
        req = rng.integers(0, 64, size=5000).astype(np.int32)
        pos = np.array([rng.integers(-1, lens[r]) for r in req], np.int32)      # -1 = padded row
        want = paging_ref.slot_mapping(bt.table, req, pos)
        got = ops.slot_mapping(dev(torch.from_numpy(bt.table)), dev(torch.from_numpy(req)), dev(torch.from_numpy(pos)))
        assert got.dtype == torch.int64 and np.array_equal(got.cpu().numpy(), want)
    assert ops.slot_mapping(dev(torch.zeros(1, 4, dtype=torch.int32)), dev(torch.zeros(0, dtype=torch.int32)),
                            dev(torch.zeros(0, dtype=torch.int32))).numel() == 0          # empty input


# ------------------------------------------------------------------ norms
@pytest.mark.parametrize("T", [1, 2, 7, 256])
def test_rmsnorm_and_fused_add(ops, T):
    g = torch.Generator().manual_seed(T)
    x = (torch.randn(T, 1024, generator=g) * 0.7).to(BF16)
    res = (torch.randn(T, 1024, generator=g) * 1.3).to(BF16)
    w = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    out = ops.rmsnorm(dev(x), dev(w))
    assert_close_bf16(out, rmsnorm(x.float(), w, 1e-5, "bf16"), ulps=1.01, what="rmsnorm")
    r = dev(res.clone())
    out = ops.rmsnorm(dev(x), dev(w), residual=r)
    z = (x.float() + res.float()).to(BF16)
    assert torch.equal(r.cpu(), z), "residual stream must be bf16(x + res) exactly"
    assert_close_bf16(out, rmsnorm(z.float(), w, 1e-5, "bf16"), ulps=1.01, what="add_rmsnorm")


# ------------------------------------------------------------------ linear layers
def _ref_linear(x, W):
    return (x.double() @ W.double().T).float()


@pytest.mark.parametrize("T", [1, 2, 3, 4])
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 4096), (8194, 1024)])
def test_linear_gemv(ops, T, N, K):
    g = torch.Generator().manual_seed(N + T)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(N, K, generator=g) * 0.05).to(BF16)
    y = ops.linear(dev(x), dev(W), path=1)
    # fp32 accumulation of K products, then one bf16 rounding: 1 ulp + accumulation noise
    assert_close_bf16(y, _ref_linear(x, W), ulps=1.0, atol=2e-3, what=f"gemv T={T} N={N} K={K}")


@pytest.mark.parametrize("T", [1, 5, 64, 128, 200, 256, 300, 2500, 4200])     # 2500 rows: the two-CTAs-per-SM configuration; 4200: CTA pairs (cta_group::2)
@pytest.mark.parametrize("N,K", [(3072, 1024), (1024, 1024), (1024, 4096), (8194, 1024)])
def test_linear_gemm_tcgen05(ops, T, N, K):
    g = torch.Generator().manual_seed(N + 7 * T)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(N, K, generator=g) * 0.05).to(BF16)
    ld = 8256 if N == 8194 else N
    out = torch.full((T, ld), 7.0, dtype=BF16).cuda()
    ops.linear(dev(x), dev(W), path=2, out=out)
    assert_close_bf16(out[:, :N], _ref_linear(x, W), ulps=1.0, atol=2e-3, what=f"gemm T={T} N={N} K={K}")
    if ld > N:
        assert (out[:, N:].float() == 7.0).all(), "columns >= N must not be written"


@pytest.mark.parametrize("T,path", [(2, 1), (4, 1), (6, 2), (256, 2)])
def test_linear_silu_mul_epilogue(ops, T, path):
    g = torch.Generator().manual_seed(T)
    x = torch.randn(T, 1024, generator=g).to(BF16)
    Wgu = (torch.randn(8192, 1024, generator=g) * 0.05).to(BF16)
    packed = ops.pack_gate_up(dev(Wgu))
    # packing is a pure permutation: rows [g0..15,u0..15,g16..31,...]
    j = torch.arange(4096)
    assert torch.equal(packed.cpu()[(j // 16) * 32 + j % 16], Wgu[:4096])
    assert torch.equal(packed.cpu()[(j // 16) * 32 + 16 + j % 16], Wgu[4096:])
    y = ops.linear(dev(x), packed, epilogue=1, path=path)
    gu = _ref_linear(x, Wgu).to(BF16).float()
    want = torch.nn.functional.silu(gu[:, :4096]) * gu[:, 4096:]
    # g,u may each differ by 1 ulp before the activation -> 3 ulps on the product
    assert_close_bf16(y, want, ulps=3.0, atol=3e-3, what=f"silu_mul T={T}")


# ------------------------------------------------------------------ RoPE + KV append + attention
def _build_paged(ops, cfg, lens, seed, num_blocks=None):
    """Random per-sequence K/V histories written through rope_kv_append into a paged cache."""
    S = len(lens)
    g = torch.Generator().manual_seed(seed)
    max_len = max(lens)
    nb = num_blocks or (sum(-(-n // 16) for n in lens) + 8)
    pool = paging_ref.BlockPool(nb)
    bt = paging_ref.BlockTable(pool, S, ((max_len + 15) // 16) * 16)
    order = np.random.default_rng(seed).permutation(S)
    for r in order:                                       # interleaved allocation -> scattered pages
        bt.ensure(int(r), lens[r])
    kv = torch.zeros(nb * 65536, dtype=torch.uint8).cuda()
    cs = rope_cos_sin(cfg, max_len + 1)
    cs_d = dev(cs.to(BF16))
    bt_d = dev(torch.from_numpy(bt.table))
    qkv_all = [(torch.randn(n, 3072, generator=g)).to(BF16) for n in lens]
    for r in range(S):
        n = lens[r]
        pos = torch.arange(n, dtype=torch.int32)
        slot = ops.slot_mapping(bt_d, dev(torch.full((n,), r, dtype=torch.int32)), dev(pos))
        ops.rope_kv_append(dev(qkv_all[r]), dev(pos), slot, cs_d, kv)
    return kv, bt_d, qkv_all, cs


def _ref_attn(cfg, qkv_hist, q_row, L, cs):
    """fp32 attention of one query (already rotated) over the first L tokens of a history."""
    k = apply_rope(qkv_hist[:L, 1024:2048].float().view(L, 16, 64), torch.arange(L), cs, "bf16")
    v = qkv_hist[:L, 2048:].float().view(L, 16, 64)
    s = torch.einsum("hd,lhd->hl", q_row.view(16, 64), k) * 0.125
    return torch.einsum("hl,lhd->hd", torch.softmax(s, -1), v).reshape(1024)


def test_rope_q_in_place_and_kv_layout(ops):
    cfg = T3ModelConfig()
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, [37], seed=1)
    n = 37
    qkv = dev(qkv_all[0].clone())
    pos = torch.arange(n, dtype=torch.int32)
    slot = ops.slot_mapping(bt_d, dev(torch.zeros(n, dtype=torch.int32)), dev(pos))
    ops.rope_kv_append(qkv, dev(pos), slot, dev(cs.to(BF16)), kv)
    want_q = apply_rope(qkv_all[0][:, :1024].float().view(n, 16, 64), pos.long(), cs, "bf16").reshape(n, 1024)
    assert_close_bf16(qkv[:, :1024], want_q, ulps=1.01, what="rotated q")
    assert torch.equal(qkv[:, 2048:].cpu(), qkv_all[0][:, 2048:]), "v columns untouched"
    # page layout: K[head][tok][chunk ^ (tok&7)], V[head][tok][chunk]
    want_k = apply_rope(qkv_all[0][:, 1024:2048].float().view(n, 16, 64), pos.long(), cs, "bf16").to(BF16)
    pages = kv.cpu().view(-1, 65536)
    bt = bt_d.cpu().numpy()
    for t in (0, 5, 16, 36):
        page = pages[bt[0, t // 16]].view(torch.bfloat16)
        for h in (0, 7, 15):
            krow = page[(h * 16 + t % 16) * 64:(h * 16 + t % 16) * 64 + 64].view(8, 8)
            unsw = torch.stack([krow[c ^ (t % 16 & 7)] for c in range(8)]).reshape(64)
            assert_close_bf16(unsw, want_k[t, h], ulps=1.01, what="K page")
            vrow = page[16384 + (h * 16 + t % 16) * 64:16384 + (h * 16 + t % 16) * 64 + 64]
            assert torch.equal(vrow, qkv_all[0][t, 2048 + h * 64:2048 + h * 64 + 64])
    # slot -1 rows must not touch the cache
    before = kv.clone()
    ops.rope_kv_append(dev(qkv_all[0][:3].clone()), dev(pos[:3]), dev(torch.full((3,), -1, dtype=torch.int64)),
                       dev(cs.to(BF16)), kv)
    assert torch.equal(before, kv)


@pytest.mark.parametrize("lens", [[1], [15, 16, 17], [100, 3, 450, 64], [1000, 999], list(range(1, 70))])
def test_paged_attention_vs_oracle(ops, lens):
    cfg = T3ModelConfig()
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=len(lens))
    S = len(lens)
    g = torch.Generator().manual_seed(99)
    q = torch.randn(S, 3072, generator=g).to(BF16)
    out = ops.paged_attn_decode(dev(q), kv, bt_d, dev(torch.arange(S, dtype=torch.int32)),
                                dev(torch.tensor(lens, dtype=torch.int32)), max(lens))
    for r in range(S):
        want = _ref_attn(cfg, qkv_all[r], q[r, :1024].float(), lens[r], cs)
        # fp32 softmax/accumulate, one bf16 rounding of the output
        assert_close_bf16(out[r], want, ulps=1.0, atol=2e-3, what=f"attn row {r} L={lens[r]}")


def test_paged_attention_many_rows_single_split_and_empty_rows(ops):
    """Prefill-style use: many query rows, each attending to a prefix of a shared sequence;
    kv_len = 0 rows (finished pairs) give zeros."""
    cfg = T3ModelConfig()
    lens = [200, 77]
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=3)
    rows = [(0, L) for L in range(1, 201, 1)] + [(1, L) for L in range(0, 78)] * 12   # > 148*6 rows -> nsplit 1
    g = torch.Generator().manual_seed(5)
    q = torch.randn(len(rows), 3072, generator=g).to(BF16)
    out = ops.paged_attn_decode(dev(q), kv, bt_d, dev(torch.tensor([r for r, _ in rows], dtype=torch.int32)),
                                dev(torch.tensor([L for _, L in rows], dtype=torch.int32)), 200)
    for i in list(range(0, len(rows), 37)) + [200, 201]:
        r, L = rows[i]
        want = _ref_attn(cfg, qkv_all[r], q[i, :1024].float(), L, cs) if L else torch.zeros(1024)
        assert_close_bf16(out[i], want, ulps=1.0, atol=2e-3, what=f"row {i} (seq {r}, L={L})")


def test_paged_attention_split_tickets_reset(ops):
    """Split-K rows are merged by the last-arriving CTA; the ticket counters must return to zero so
    a second launch over the same workspace gives identical bits."""
    cfg = T3ModelConfig()
    lens = [700, 650, 33, 512]
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=8)
    g = torch.Generator().manual_seed(4)
    q = dev(torch.randn(4, 3072, generator=g).to(BF16))
    args = (q, kv, bt_d, dev(torch.arange(4, dtype=torch.int32)), dev(torch.tensor(lens, dtype=torch.int32)), 700)
    assert ops._lib.load().cbx_attn_num_splits(4, 700) > 1
    a = ops.paged_attn_decode(*args).clone()
    for _ in range(3):
        assert torch.equal(ops.paged_attn_decode(*args), a)
    for r in range(4):
        want = _ref_attn(cfg, qkv_all[r], q[r, :1024].float().cpu(), lens[r], cs)
        assert_close_bf16(a[r], want, ulps=1.0, atol=2e-3, what=f"split attn row {r}")


@pytest.mark.parametrize("case", ["ragged", "long_few", "zeros_at_edges", "beyond_plan"])
def test_paged_attention_balanced_page_ranges(ops, case):
    """4 < T <= 4096 rows: the launch's pages are cut into 148 equal ranges, rows cut by a range
    boundary are merged by ticket.  Ragged lengths, rows spanning many CTAs, kv_len = 0 rows at the
    start / end / around range boundaries; a second launch over the same workspace must give the
    same bits (tickets self-reset).  'beyond_plan' (T > 4096) takes the fixed-split kernel."""
    cfg = T3ModelConfig()
    rng = random.Random(11)
    if case == "ragged":
        lens = [850, 333, 64]
        rows = [(rng.randrange(3), 0) for _ in range(300)]
        rows = [(s, rng.randint(1, lens[s])) for s, _ in rows]
    elif case == "long_few":
        lens = [1000, 977]
        rows = [(0, 1000), (1, 977), (0, 999), (1, 1), (0, 17), (1, 500), (0, 3)]
    elif case == "zeros_at_edges":
        lens = [400, 90]
        rows = [(0, 0)] * 3 + [(rng.randrange(2), 0) for _ in range(200)] + [(1, 0)] * 5
        rows = [(s, 0 if i < 3 or i >= 203 or i % 7 == 0 else rng.randint(1, lens[s])) for i, (s, _) in enumerate(rows)]
    else:
        lens = [40, 33]
        rows = [(i & 1, 1 + (i * 7) % lens[i & 1]) for i in range(4100)]
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=21)
    g = torch.Generator().manual_seed(6)
    q = torch.randn(len(rows), 3072, generator=g).to(BF16)
    args = (dev(q), kv, bt_d, dev(torch.tensor([r for r, _ in rows], dtype=torch.int32)),
            dev(torch.tensor([L for _, L in rows], dtype=torch.int32)), max(lens))
    out = ops.paged_attn_decode(*args).clone()
    assert torch.equal(ops.paged_attn_decode(*args), out)
    # plan built once and reused by later launches over the same kv_len (what the engine does per step)
    assert torch.equal(ops.paged_attn_decode(*args, plan="make"), out)
    assert torch.equal(ops.paged_attn_decode(*args, plan="reuse"), out)
    check = range(len(rows)) if len(rows) <= 400 else range(0, len(rows), 41)
    for i in check:
        r, L = rows[i]
        want = _ref_attn(cfg, qkv_all[r], q[i, :1024].float(), L, cs) if L else torch.zeros(1024)
        assert_close_bf16(out[i], want, ulps=1.0, atol=2e-3, what=f"{case} row {i} (seq {r}, L={L})")


# ------------------------------------------------------------------ fused layer pieces
def _unswizzled_k(page_bf16, h, tok):
    krow = page_bf16[(h * 16 + tok) * 64:(h * 16 + tok) * 64 + 64].view(8, 8)
    return torch.stack([krow[c ^ (tok & 7)] for c in range(8)]).reshape(64)


@pytest.mark.parametrize("T", [1, 2, 4, 6, 200, 1600, 8300])         # 1600 rows: 128-wide tiles, two heads per RoPE epilogue; 8300: CTA pairs, four heads
def test_layer_qkv_fused_norm_rope_append(ops, T):
    """[RMSNorm → QKV → RoPE → KV append] in one op (GEMV kernel for T <= 4, rmsnorm + tcgen05 GEMM
    with the RoPE/append epilogue above) against the oracle's op-by-op bf16 chain."""
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
        slot[T - 1] = -1                                   # finished pair: nothing may be cached
    kv = torch.zeros(n_blocks * 65536, dtype=torch.uint8).cuda()
    cs = rope_cos_sin(cfg, 1024)
    qkv = ops.layer_qkv(dev(res), dev(ln), dev(W), dev(cs.to(BF16)), dev(pos), dev(slot), kv)
    h = rmsnorm(res.float(), ln, 1e-5, "bf16")
    y = _ref_linear(h, W).to(BF16).float()
    want_q = apply_rope(y[:, :1024].view(T, 16, 64), pos.long(), cs, "bf16").reshape(T, 1024)
    want_k = apply_rope(y[:, 1024:2048].view(T, 16, 64), pos.long(), cs, "bf16").reshape(T, 1024)
    want_v = y[:, 2048:]
    # y carries <= 1 ulp of accumulation-order noise per element; rotation mixes two of them
    tol = dict(ulps=3.0, atol=0.03)
    assert_close_bf16(qkv[:, :1024], want_q, what="fused q", **tol)
    pages = kv.cpu().view(-1, 65536)
    touched = set()
    for t in range(T):
        if slot[t] < 0:
            continue
        b, tok = int(slot[t]) // 16, int(slot[t]) % 16
        touched.add(b)
        page = pages[b].view(torch.bfloat16)
        for hd in range(16):
            assert_close_bf16(_unswizzled_k(page, hd, tok), want_k[t, hd * 64:hd * 64 + 64], what="fused K", **tol)
            vrow = page[16384 + (hd * 16 + tok) * 64:16384 + (hd * 16 + tok) * 64 + 64]
            assert_close_bf16(vrow, want_v[t, hd * 64:hd * 64 + 64], ulps=1.01, atol=0.02, what="fused V")
    for b in range(n_blocks):
        if b not in touched:
            assert not pages[b].any(), f"page {b} must stay untouched"


@pytest.mark.parametrize("T", [1, 3, 4, 5, 256, 4800])
@pytest.mark.parametrize("K", [1024, 4096])
def test_linear_residual_in_place(ops, T, K):
    g = torch.Generator().manual_seed(T * 7 + K)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(1024, K, generator=g) * 0.03).to(BF16)
    res = (torch.randn(T, 1024, generator=g) * 2).to(BF16)
    got = ops.linear_residual(dev(x), dev(W), dev(res.clone()))
    y = _ref_linear(x, W).to(BF16).float()
    want = (y + res.float()).to(BF16).float()
    err = (got.float().cpu() - want).abs()
    tol = 2.0 * ULP * y.abs() + 1.01 * ULP * want.abs() + 1e-6     # 2 ulps on the GEMM term + the final rounding
    assert (err <= tol).all(), f"max err {err.max():.3e}"


@pytest.mark.parametrize("T", [1, 2, 4, 7, 256, 700])
def test_layer_gate_up_fused_norm_silu(ops, T):
    g = torch.Generator().manual_seed(T + 31)
    res = (torch.randn(T, 1024, generator=g) * 1.2).to(BF16)
    ln = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    Wgu = (torch.randn(8192, 1024, generator=g) * 0.03).to(BF16)
    act = ops.layer_gate_up(dev(res), dev(ln), ops.pack_gate_up(dev(Wgu)))
    h = rmsnorm(res.float(), ln, 1e-5, "bf16")
    gu = _ref_linear(h, Wgu).to(BF16).float()
    want = torch.nn.functional.silu(gu[:, :4096]) * gu[:, 4096:]
    assert_close_bf16(act, want, ulps=3.0, atol=3e-3, what=f"fused gate_up T={T}")


@pytest.mark.parametrize("T", [2, 4, 6, 256])
def test_head_logits_fused_final_norm(ops, T):
    g = torch.Generator().manual_seed(T + 77)
    res = (torch.randn(T, 1024, generator=g) * 1.2).to(BF16)
    nw = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    Wh = (torch.randn(8194, 1024, generator=g) * 0.05).to(BF16)
    lg = ops.head_logits(dev(res), dev(nw), dev(Wh))
    h = rmsnorm(res.float(), nw, 1e-5, "bf16")
    assert_close_bf16(lg[:, :8194], _ref_linear(h, Wh), ulps=2.0, atol=0.02, what=f"head T={T}")


@pytest.mark.parametrize("T,K,splits", [(5, 1024, 4), (256, 1024, 4), (256, 4096, 4), (300, 4096, 2), (128, 4096, 1)])
def test_split_k_partials_and_reduce_norm(ops, T, K, splits):
    """o_proj / down_proj as split-K partials folded by the following add + RMSNorm kernel: same
    rounding points as the unsplit chain (bf16(y), bf16(y + res), norm)."""
    g = torch.Generator().manual_seed(T + K + splits)
    x = torch.randn(T, K, generator=g).to(BF16)
    W = (torch.randn(1024, K, generator=g) * 0.03).to(BF16)
    res = (torch.randn(T, 1024, generator=g) * 2).to(BF16)
    w = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    part = ops.linear_partials(dev(x), dev(W), splits)
    y64 = x.double() @ W.double().T
    assert (part.double().sum(0).cpu() - y64).abs().max() < 2e-3, "partials must sum to the full product"
    r = dev(res.clone())
    out = ops.rmsnorm_reduce(part, r, dev(w))
    y = y64.float().to(BF16).float()
    want_res = (y + res.float()).to(BF16).float()
    err = (r.float().cpu() - want_res).abs()
    assert (err <= 2.0 * ULP * y.abs() + 1.01 * ULP * want_res.abs() + 1e-6).all(), f"res max err {err.max():.3e}"
    # the norm is checked on the kernel's own residual (so only the norm's rounding is in play)
    assert_close_bf16(out, rmsnorm(r.float().cpu(), w, 1e-5, "bf16"), ulps=1.01, what="norm after reduce")


@pytest.mark.parametrize("lens", [[1], [16], [17, 5], [200, 77, 33], [450]])
def test_paged_attention_prefill_causal(ops, lens):
    """Causal prefill attention (16 query rows share each page load, tensor-core QK^T / PV) against the
    fp32 reference, for every row of every sequence, and against the per-row decode kernel.
    Tolerance: like FlashAttention-2 (the kernel vLLM runs here) the probabilities are rounded to bf16
    for the PV product, so |err| <= 2^-9 · sum_j p_j |v_j| / l <= 2^-9 · max|v| in ABSOLUTE terms (it does
    not shrink when the v_j cancel): atol = 2^-9 · 4 for the N(0,1) values used here, plus 2 bf16 ulps."""
    cfg = T3ModelConfig()
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=10 + len(lens))
    starts = np.concatenate(([0], np.cumsum(lens)[:-1])).tolist()
    T = sum(lens)
    g = torch.Generator().manual_seed(7)
    q = torch.randn(T, 3072, generator=g).to(BF16)
    items = ops.prefill_items(starts, lens, list(range(len(lens)))).cuda()
    out = ops.paged_attn_prefill(dev(q), kv, bt_d, items)
    rows = [(s, p) for s, n in enumerate(lens) for p in range(n)]
    dec = ops.paged_attn_decode(dev(q), kv, bt_d, dev(torch.tensor([s for s, _ in rows], dtype=torch.int32)),
                                dev(torch.tensor([p + 1 for _, p in rows], dtype=torch.int32)), max(lens))
    assert_close_bf16(out, dec, ulps=2.0, atol=8e-3, what="prefill vs per-row decode kernel")
    for i in list(range(0, T, max(1, T // 40))) + [T - 1]:
        s, p = rows[i]
        want = _ref_attn(cfg, qkv_all[s], q[i, :1024].float(), p + 1, cs)
        assert_close_bf16(out[i], want, ulps=2.0, atol=8e-3, what=f"prefill row {i} (seq {s}, pos {p})")


# ------------------------------------------------------------------ token post-processing
def test_postprocess_tokens_matches_oracle_bit_exact():
    """cbx_postprocess_tokens against the CPU restatement of tts.py:448-465 + drop_invalid_tokens
    on adversarial streams: stop early / late / absent, SOS and EOS inside, invalid ids, empty
    results, guard caps, slot -> row scatter."""
    import numpy as np
    from chatterbox_vllm_b200 import ops
    from oracle import prompt_ref
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(7)
    B, stride = 24, 700
    sb = ops.SamplerBuffers(B, stride, dev)
    toks = rng.integers(0, 6561, size=(B, stride)).astype(np.int32)
    n_out = rng.integers(0, stride + 1, size=B).astype(np.int32)
    n_out[0], n_out[1], n_out[2] = 0, 1, stride
    for b in range(B):
        n = int(n_out[b])
        if n == 0:
            continue
        if b % 3 != 0:                                   # a stop token somewhere (not always the last)
            toks[b, rng.integers(0, n)] = 6562
        if b % 4 == 1:
            toks[b, rng.integers(0, n)] = 6561           # an SOS inside
        if b % 5 == 2:
            toks[b, rng.integers(0, n, size=5)] = rng.integers(6563, 8194, size=5)   # ids beyond the speech vocab
    toks[3, 0] = 6562                                    # stop first: empty result
    toks[4, :2] = (5, 6562)                              # stop inside the cropped tail
    sb.out_tokens.copy_(torch.from_numpy(toks))
    sb.n_out.copy_(torch.from_numpy(n_out))
    for crop, use_caps in ((2, False), (0, True), (5, True)):
        slots = torch.tensor(rng.permutation(B).astype(np.int32), device=dev)
        rows = torch.tensor(rng.permutation(B).astype(np.int32), device=dev)
        caps = rng.integers(1, 400, size=B).astype(np.int32) if use_caps else None
        dst = torch.full((B, stride), -7, dtype=torch.int64, device=dev)
        dlen = torch.full((B,), -1, dtype=torch.int32, device=dev)
        ops.postprocess_tokens(sb, slots, dst, rows, dlen, crop, None if caps is None else torch.from_numpy(caps).to(dev))
        dst, dlen = dst.cpu().numpy(), dlen.cpu().numpy()
        for i in range(B):
            b, r = int(slots[i]), int(rows[i])
            want = prompt_ref.postprocess_tokens((toks[b, :n_out[b]].astype(np.int64) + 1000).tolist(), crop)
            if caps is not None:
                want = want[:caps[i]]
            assert dlen[r] == len(want), f"slot {b}: {dlen[r]} vs {len(want)} tokens (crop {crop})"
            assert np.array_equal(dst[r, :len(want)], want), f"slot {b} (crop {crop})"
            assert (dst[r, len(want):] == -7).all(), "nothing may be written past the result"


# ------------------------------------------------------------------ conditioning encoder
COND_REL_TOL = 2e-5      # fp32 kernels vs the fp32 reference modules: summation order only


def test_cond_encoder_matches_reference_modules_fp32():
    """cbx_cond_encode (T3CondEnc + Perceiver, fp32 FMA) against tests/golden/cond_enc.npz — the output
    of the reference's own modules — and against the oracle on a ragged prompt length."""
    import os
    import numpy as np
    from chatterbox_vllm_b200.cond_enc import T3CondEncoder
    from oracle import cond_ref
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cond_enc.npz"))
    for c in (0, 1):
        wseed, iseed, n = g[f"meta{c}"].tolist()
        sd = cond_ref.random_state_dict(wseed)
        speech_emb, speech_pos, spk, tok = cond_ref.random_inputs(iseed, n)
        enc = T3CondEncoder({"cond_enc." + k: v for k, v in sd.items()}, speech_emb, speech_pos)
        got = enc(spk, tok, float(g[f"emotion{c}"][0]))
        want = torch.from_numpy(g[f"cond{c}"])
        assert got.is_cuda and got.dtype == torch.float32 and got.shape == (34, 1024)
        scale = float(want.abs().max())
        err = float((got.cpu() - want).abs().max())
        assert err <= COND_REL_TOL * scale, f"case {c}: max-abs {err:.3e} at scale {scale:.2f}"
        assert torch.equal(got[33].cpu(), want[33]), "the emotion row is one multiply: bit-exact"
        row = enc.emotion_row(0.3)
        assert torch.equal(row.cpu(), torch.from_numpy(g[f"row03_{c}"]))
        # another prompt length through the same encoder object (workspace reuse), vs the oracle
        for n2 in (1, 33, 149):
            got2 = enc(spk, tok[:n2], 0.5).cpu()
            want2 = cond_ref.cond_encode(sd, spk, tok[:n2], speech_emb, speech_pos, 0.5)
            assert float((got2 - want2).abs().max()) <= COND_REL_TOL * float(want2.abs().max()), f"n_prompt={n2}"
    with pytest.raises(ValueError):
        enc(spk, tok[:0])


def test_paged_attention_prefill_is_deterministic(ops):
    """Regression for a write-after-read race in the prefill attention ring: ptxas scheduled the stage-release
    mbarrier arrive above the MMAs consuming the last ldmatrix, so the TMA refill of a stage could overtake a
    shared-memory read still in flight — a few heads of a few 16-row items then saw a neighbouring page, differently
    on every launch (inside the parity tolerance at test sizes; found by running BASELINE config 2 twice at full
    size).  Fixed with a proxy fence before the release; identical inputs must now give identical bits."""
    torch.manual_seed(0)
    S, P = 32, 200
    nblk = -(-P // 16)
    kv = (torch.randn(((S * nblk + 1) * 65536) // 2, device="cuda") * 0.5).to(BF16).view(torch.uint8)
    bt = torch.zeros(S, 64, dtype=torch.int32, device="cuda")
    bt[:, :nblk] = (torch.randperm(S * nblk, device="cuda").to(torch.int32) + 1).view(S, nblk)
    qkv = torch.randn(S * P, 3072, device="cuda").to(BF16)
    items = ops.prefill_items([i * P for i in range(S)], [P] * S, list(range(S))).cuda()
    out = torch.zeros(S * P, 1024, dtype=BF16, device="cuda")
    ref = None
    for _ in range(12):
        ops.paged_attn_prefill(qkv, kv, bt, items, out=out)
        cur = out.clone()
        if ref is None:
            ref = cur
        assert torch.equal(cur, ref), "prefill attention: same inputs, different bits"


def test_prefill_rows_equal_the_row_by_row_builder(ops):
    """cbx_prefill_rows against the numpy builder T3Engine.prefill used before it (the reference's per-request Python of
    T3MultiModalProcessor.apply / get_input_embeddings, t3.py:229-262,548-569, for both members of every pair): ragged
    text lengths, histories of preempted requests, scattered block-table rows.  Integer work: bit-exact."""
    rng = np.random.default_rng(5)
    B = 37
    seq, text_flat, hist_flat, want_desc, want_pos, want_rseq = [], [], [], [], [], []
    T = t_off = h_off = 0
    for j in range(9):
        nt = int(rng.integers(1, 60))
        nh = int(rng.integers(0, 40)) if j % 3 == 0 else 0
        ids = rng.integers(0, 700, nt).astype(np.int32)
        hist = rng.integers(0, 6561, nh).astype(np.int32)
        P = 34 + nt + 1 + nh
        slot = int(rng.integers(0, B))
        text_flat += ids.tolist()
        hist_flat += hist.tolist()
        cond_rows = np.zeros((34, 4), np.int32)
        cond_rows[:, 1], cond_rows[:, 2] = j, np.arange(34)
        text_c = np.zeros((nt, 4), np.int32)
        text_c[:, 0], text_c[:, 1], text_c[:, 2] = 1, ids, np.arange(nt)
        text_u = np.zeros((nt, 4), np.int32)
        text_u[:, 0] = 2
        hrows = np.zeros((nh, 4), np.int32)
        hrows[:, 0], hrows[:, 1] = 4, hist
        sos = np.array([[3, 6561, 0, 0]], np.int32)
        for half, text in ((0, text_c), (1, text_u)):
            seq.append((T, P, nt, j, half, half * B + slot, t_off, h_off))
            want_desc += [cond_rows, text, sos, hrows]
            want_pos.append(np.arange(P, dtype=np.int32))
            want_rseq.append(np.full(P, half * B + slot, np.int32))
            T += P
        t_off += nt
        h_off += nh
    i32 = lambda v: torch.tensor(v if len(v) else [0], dtype=torch.int32).cuda()  # noqa: E731
    desc, pos, rseq, kvl = ops.prefill_rows(i32(seq).view(-1, 8), i32(text_flat), i32(hist_flat), T)
    assert np.array_equal(desc.cpu().numpy(), np.concatenate(want_desc))
    assert np.array_equal(pos.cpu().numpy(), np.concatenate(want_pos))
    assert np.array_equal(rseq.cpu().numpy(), np.concatenate(want_rseq))
    assert np.array_equal(kvl.cpu().numpy(), np.concatenate(want_pos) + 1)
