"""cbx_paged_attn_decode_static — the allocation-free decode attention entry point (scratch in a
file-scope device array inside libcbx_t3.so) — against the existing cbx_paged_attn_decode, through
the C-ABI, on two calls in a row.  Collected last on purpose; it
builds everything it needs itself, so it passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest
import torch

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


def _paged_case(ops, lens, n_rows, seed):
    """Random K/V histories appended through the library's own RoPE + KV-append into scattered pages,
    and n_rows query rows with ragged kv_len (some zero) over them."""
    g = torch.Generator().manual_seed(seed)
    rng = np.random.default_rng(seed)
    S, max_len = len(lens), max(lens)
    mb = (max_len + 15) // 16
    need = [(n + 15) // 16 for n in lens]
    ids = rng.permutation(np.arange(1, sum(need) + 9))            # block 0 is the null block
    bt = np.zeros((S, mb), np.int32)
    k = 0
    for r in range(S):
        bt[r, :need[r]] = ids[k:k + need[r]]
        k += need[r]
    bt_d = torch.from_numpy(bt).cuda()
    kv = torch.zeros((sum(need) + 9) * 65536, dtype=torch.uint8, device="cuda")
    inv = 1.0 / (10000.0 ** (torch.arange(0, 64, 2).float() / 64))
    ang = torch.arange(max_len + 1).float()[:, None] * inv[None, :]
    cs = torch.cat([ang.cos(), ang.sin()], -1).to(BF16).cuda()     # any table does: both entry points read the same cache
    for r in range(S):
        n = lens[r]
        pos = torch.arange(n, dtype=torch.int32).cuda()
        slot = ops.slot_mapping(bt_d, torch.full((n,), r, dtype=torch.int32).cuda(), pos)
        ops.rope_kv_append(torch.randn(n, 3072, generator=g).to(BF16).cuda(), pos, slot, cs, kv)
    row_seq = torch.from_numpy(rng.integers(0, S, n_rows).astype(np.int32))
    kv_len = torch.tensor([0 if i % 11 == 3 else int(rng.integers(1, lens[int(s)] + 1)) for i, s in enumerate(row_seq)],
                          dtype=torch.int32)
    q = torch.randn(n_rows, 3072, generator=g).to(BF16)
    return q.cuda(), kv, bt_d, row_seq.cuda(), kv_len.cuda(), max_len


def _close(got, want, what):
    # the bar of tests/test_gpu_kernels.py's attention tests: 1 bf16 ulp + 2e-3
    got, want = got.float().cpu(), want.float().cpu()
    err = (got - want).abs()
    bad = err > 1.0 * ULP * want.abs() + 2e-3
    assert not bad.any(), f"{what}: {int(bad.sum())} elements off, max err {err.max():.3e}"


def test_static_scratch_attention_matches_existing_entry_point_twice(ops):
    cases = [_paged_case(ops, [850, 333, 64, 500], 300, seed=5), _paged_case(ops, [120, 1000], 37, seed=6)]
    want = [ops.paged_attn_decode(*c).clone() for c in cases]
    torch.cuda.synchronize()
    outs = [torch.full((c[0].shape[0], 1024), float("nan"), dtype=BF16, device="cuda") for c in cases]
    # two calls in a row over the SAME library-held scratch, different shapes: the second must not see the first
    ops.paged_attn_decode_static(*cases[0], outs[0])
    ops.paged_attn_decode_static(*cases[1], outs[1])
    torch.cuda.synchronize()
    for i in range(2):
        _close(outs[i], want[i], f"static call {i}")
        assert torch.equal(outs[i], want[i]), f"static call {i}: same kernel, same plan -> same bits"
    # and again in the other order, with no allocation in between: free device memory must not move
    free0 = torch.cuda.mem_get_info()[0]
    for o in outs:
        o.fill_(float("nan"))
    ops.paged_attn_decode_static(*cases[1], outs[1])
    ops.paged_attn_decode_static(*cases[0], outs[0])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static entry point must not allocate device memory"
    for i in range(2):
        _close(outs[i], want[i], f"static repeat {i}")


def test_static_scratch_attention_rejects_sizes_beyond_its_array(ops):
    q, kv, bt, row_seq, kv_len, max_len = _paged_case(ops, [40, 33], 8, seed=7)
    for T in (4096, 3):
        qq = torch.zeros(T, 3072, dtype=BF16, device="cuda")
        rs = torch.zeros(T, dtype=torch.int32, device="cuda")
        kl = torch.ones(T, dtype=torch.int32, device="cuda")
        out = torch.zeros(T, 1024, dtype=BF16, device="cuda")
        with pytest.raises(RuntimeError, match="outside 5..4095"):
            ops.paged_attn_decode_static(qq, kv, bt, rs, kl, max_len, out)
    out = torch.zeros(8, 1024, dtype=BF16, device="cuda")
    ops.paged_attn_decode_static(q, kv, bt, row_seq, kv_len, max_len, out)         # still usable after a rejection
    _close(out, ops.paged_attn_decode(q, kv, bt, row_seq, kv_len, max_len), "after rejection")
