"""cbx_layer_attn_block — the attention half of a layer as one cluster kernel (k_block.cu, T <= 4 rows) — through
the C-ABI against (a) the three-kernel chain it replaces (layer_qkv -> paged_attn_decode -> linear_residual) and
(b) the oracle's op-by-op bf16 chain in fp32 torch.  The new token's K/V and the rotated q use the GEMV kernel's
summation order, so they are required bit for bit; the residual differs from the chain by summation order in the
attention merge and the O-projection only."""
import numpy as np
import pytest
import torch

from oracle import paging_ref
from oracle.t3_ref import T3ModelConfig, apply_rope, rmsnorm, rope_cos_sin

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


def _case(ops, hist, active, seed):
    """hist[r] = tokens already cached for row r's sequence; the step appends one token per ACTIVE row."""
    cfg = T3ModelConfig()
    T = len(hist)
    g = torch.Generator().manual_seed(seed)
    max_len = max(hist) + 1
    nb = sum(-(-(n + 1) // 16) for n in hist) + 8
    pool = paging_ref.BlockPool(nb)
    bt = paging_ref.BlockTable(pool, T, ((max_len + 15) // 16) * 16)
    for r in np.random.default_rng(seed).permutation(T):
        bt.ensure(int(r), hist[r] + 1)
    kv = torch.zeros(nb * 65536, dtype=torch.uint8).cuda()
    cs = rope_cos_sin(cfg, max_len + 1)
    cs_d, bt_d = cs.to(BF16).cuda(), torch.from_numpy(bt.table).cuda()
    past = [torch.randn(n, 3072, generator=g).to(BF16) for n in hist]
    for r, n in enumerate(hist):
        if n:
            p = torch.arange(n, dtype=torch.int32).cuda()
            ops.rope_kv_append(past[r].cuda(), p, ops.slot_mapping(bt_d, torch.full((n,), r, dtype=torch.int32).cuda(), p), cs_d, kv)
    res = (torch.randn(T, 1024, generator=g) * 1.5).to(BF16)
    ln = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16)
    wqkv = (torch.randn(3072, 1024, generator=g) * 0.03).to(BF16)
    wo = (torch.randn(1024, 1024, generator=g) * 0.03).to(BF16)
    pos = torch.tensor(hist, dtype=torch.int32)
    row_seq = torch.arange(T, dtype=torch.int32)
    act = torch.tensor(active)
    kv_len = torch.where(act, pos + 1, torch.zeros_like(pos))
    slot = ops.slot_mapping(bt_d, row_seq.cuda(), pos.cuda()).cpu()
    slot = torch.where(act, slot, torch.full_like(slot, -1))
    return dict(cfg=cfg, T=T, kv=kv, cs=cs, cs_d=cs_d, bt=bt_d, past=past, res=res, ln=ln, wqkv=wqkv, wo=wo, pos=pos,
                row_seq=row_seq, kv_len=kv_len, slot=slot, max_len=max_len, active=active, hist=hist)


def _chain(ops, c, kv):
    d = lambda t: t.cuda()  # noqa: E731
    qkv = ops.layer_qkv(d(c["res"]), d(c["ln"]), d(c["wqkv"]), c["cs_d"], d(c["pos"]), d(c["slot"]), kv)
    attn = ops.paged_attn_decode(qkv, kv, c["bt"], d(c["row_seq"]), d(c["kv_len"]), c["max_len"])
    return qkv, ops.linear_residual(attn, d(c["wo"]), d(c["res"].clone()))


def _fused(ops, c, kv, q_out=None):
    d = lambda t: t.cuda()  # noqa: E731
    return ops.layer_attn_block(d(c["res"].clone()), d(c["ln"]), d(c["wqkv"]), d(c["wo"]), c["cs_d"], d(c["pos"]), d(c["slot"]),
                                kv, c["bt"], d(c["row_seq"]), d(c["kv_len"]), c["max_len"], q_out=q_out)


def _reference(c):
    """fp32 torch, bf16 rounding where the chain materialises a bf16 tensor."""
    r16 = lambda x: x.to(BF16).float()  # noqa: E731
    T, cs = c["T"], c["cs"]
    h = rmsnorm(c["res"].float(), c["ln"], 1e-5, "bf16")
    y = r16(h @ c["wqkv"].float().T)
    q = apply_rope(y[:, :1024].view(T, 16, 64), c["pos"].long(), cs, "bf16")
    k = apply_rope(y[:, 1024:2048].view(T, 16, 64), c["pos"].long(), cs, "bf16")
    v = y[:, 2048:].view(T, 16, 64)
    attn = torch.zeros(T, 1024)
    for t in range(T):
        if not c["active"][t]:
            continue
        n = c["hist"][t]
        pk = apply_rope(c["past"][t][:, 1024:2048].float().view(n, 16, 64), torch.arange(n), cs, "bf16")
        K = torch.cat([pk, k[t:t + 1]], 0)
        V = torch.cat([c["past"][t][:, 2048:].float().view(n, 16, 64), v[t:t + 1]], 0)
        s = torch.einsum("hd,lhd->hl", q[t], K) * 0.125
        attn[t] = torch.einsum("hl,lhd->hd", torch.softmax(s, -1), V).reshape(1024)
    o = r16(r16(attn) @ c["wo"].float().T)
    return r16(o + c["res"].float()), o


CASES = {
    "b1_ctx335": ([335, 335], [True, True]),
    "b1_fresh_and_short": ([0, 17], [True, True]),
    "one_row_long": ([3000], [True]),
    "two_rows_page_edges": ([16, 1919], [True, True]),
    "b2_ragged": ([959, 950, 15, 31], [True, True, True, True]),
    "three_rows": ([200, 77, 450], [True, True, True]),
    "finished_pair_rows": ([120, 64, 300, 301], [True, False, True, False]),
}


@pytest.mark.parametrize("name", list(CASES))
def test_fused_attention_block_vs_chain_and_reference(ops, name):
    hist, active = CASES[name]
    c = _case(ops, hist, active, seed=len(name) + sum(hist))
    assert ops._lib.load().cbx_layer_attn_block_supported(c["T"], c["max_len"]) == 1
    kv_chain, kv_fused = c["kv"].clone(), c["kv"].clone()
    qkv, res_chain = _chain(ops, c, kv_chain)
    q_out = torch.zeros(c["T"], 3072, dtype=BF16, device="cuda")
    res_fused = _fused(ops, c, kv_fused, q_out).clone()
    torch.cuda.synchronize()
    # integer-exact pieces: the appended K/V rows and the rotated q carry the GEMV kernel's bits
    assert torch.equal(kv_fused, kv_chain), "KV pages after the append differ from the chain's"
    assert torch.equal(q_out[:, :1024], qkv[:, :1024]), "rotated q differs from the chain's"
    want, o = _reference(c)
    for got, what in ((res_fused, "fused"), (res_chain, "chain")):
        err = (got.float().cpu() - want).abs()
        # 3 ulps on the o-proj term (attention output is itself 1 bf16 ulp loose) + the final rounding
        tol = 3.0 * ULP * o.abs() + 1.01 * ULP * want.abs() + 2e-3
        assert (err <= tol).all(), f"{name} {what}: max err {err.max():.3e}, {int((err > tol).sum())} elements off"
    inactive = [t for t in range(c["T"]) if not active[t]]
    for t in inactive:
        assert torch.equal(res_fused[t].cpu(), c["res"][t]), "inactive row must keep its residual"
    # the tickets reset themselves: a second launch over the same scratch gives the same bits
    kv2 = c["kv"].clone()
    assert torch.equal(_fused(ops, c, kv2), res_fused) and torch.equal(kv2, kv_fused)


def test_fused_attention_block_context_limits(ops):
    lib = ops._lib.load()
    assert lib.cbx_layer_attn_block_supported(1, 3841) == 1 and lib.cbx_layer_attn_block_supported(1, 3842) == 0
    assert lib.cbx_layer_attn_block_supported(2, 1921) == 1 and lib.cbx_layer_attn_block_supported(2, 1922) == 0
    assert lib.cbx_layer_attn_block_supported(4, 961) == 1 and lib.cbx_layer_attn_block_supported(3, 962) == 0
    assert lib.cbx_layer_attn_block_supported(5, 100) == 0
    c = _case(ops, [10, 10], [True, True], seed=3)
    c["max_len"] = 5000
    with pytest.raises(RuntimeError, match="exceeds the fused block"):
        _fused(ops, c, c["kv"].clone())
