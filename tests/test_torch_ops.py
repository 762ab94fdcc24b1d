"""The C-ABI entry points as torch.library custom ops (chatterbox_vllm_b200/torch_ops.py): schemas, fake-tensor
propagation and the no-CPU-fallback rule on CPU; on the GPU, torch.ops.cbx.* must give the bits of the ctypes path."""
import pytest
import torch

import chatterbox_vllm_b200.torch_ops as T

BF16 = torch.bfloat16


def test_ops_are_registered_with_schemas_and_mutation_annotations():
    for name in T.OPS:
        op = getattr(torch.ops.cbx, name).default
        assert str(op._schema).startswith(f"cbx::{name}(")
    s = str(torch.ops.cbx.linear_residual.default._schema)
    assert "Tensor(a2!) res" in s and s.endswith("-> ()")                      # in-place residual update is declared
    s = str(torch.ops.cbx.layer_attn_block.default._schema)
    assert "Tensor(a0!) res" in s and "Tensor(a7!) kv_layer" in s
    assert "Tensor(a6!) kv_layer" in str(torch.ops.cbx.layer_qkv.default._schema)
    assert "Tensor(a0!) res" in str(torch.ops.cbx.layer_mlp_block.default._schema)


def test_fake_tensor_propagation_needs_no_gpu():
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        d = dict(device="cuda")
        q = torch.empty(7, 3072, dtype=BF16, **d)
        kv = torch.empty(4 * 65536, dtype=torch.uint8, **d)
        bt = torch.empty(7, 8, dtype=torch.int32, **d)
        i32 = torch.empty(7, dtype=torch.int32, **d)
        res, w = torch.empty(7, 1024, dtype=BF16, **d), torch.empty(1024, dtype=BF16, **d)
        out = torch.ops.cbx.paged_attn_decode(q, kv, bt, i32, i32, 100)
        assert out.shape == (7, 1024) and out.dtype == BF16 and out.device.type == "cuda"
        assert torch.ops.cbx.slot_mapping(bt, i32, i32).dtype == torch.int64
        assert torch.ops.cbx.rmsnorm(res, w).shape == (7, 1024)
        assert torch.ops.cbx.linear(res, torch.empty(8192, 1024, dtype=BF16, **d), 1).shape == (7, 4096)
        assert torch.ops.cbx.linear(res, torch.empty(3072, 1024, dtype=BF16, **d)).shape == (7, 3072)
        assert torch.ops.cbx.layer_qkv(res, w, torch.empty(3072, 1024, dtype=BF16, **d), torch.empty(2048, 64, dtype=BF16, **d),
                                       i32, torch.empty(7, dtype=torch.int64, **d), kv).shape == (7, 3072)
        assert torch.ops.cbx.layer_gate_up(res, w, torch.empty(8192, 1024, dtype=BF16, **d)).shape == (7, 4096)
        assert torch.ops.cbx.head_logits(res, w, torch.empty(8194, 1024, dtype=BF16, **d)).shape == (7, 8256)
        assert torch.ops.cbx.linear_residual(res, torch.empty(1024, 1024, dtype=BF16, **d), res) is None


def test_no_cpu_kernel_behind_the_ops():
    with pytest.raises(NotImplementedError):
        torch.ops.cbx.rmsnorm(torch.zeros(2, 1024, dtype=BF16), torch.ones(1024, dtype=BF16))


@pytest.mark.gpu
def test_torch_ops_give_the_ctypes_paths_bits():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from chatterbox_vllm_b200 import ops
    g = torch.Generator().manual_seed(0)
    T_, n_blk = 9, 12
    res = (torch.randn(T_, 1024, generator=g) * 1.5).to(BF16).cuda()
    ln = (1 + 0.1 * torch.randn(1024, generator=g)).to(BF16).cuda()
    wqkv = (torch.randn(3072, 1024, generator=g) * 0.03).to(BF16).cuda()
    wo = (torch.randn(1024, 1024, generator=g) * 0.03).to(BF16).cuda()
    inv = 1.0 / (10000.0 ** (torch.arange(0, 64, 2).float() / 64))
    ang = torch.arange(64).float()[:, None] * inv[None, :]
    cs = torch.cat([ang.cos(), ang.sin()], -1).to(BF16).cuda()
    bt = torch.arange(1, T_ + 1, dtype=torch.int32).view(T_, 1).cuda()
    rows = torch.arange(T_, dtype=torch.int32).cuda()
    pos = torch.full((T_,), 5, dtype=torch.int32).cuda()
    kv_len = pos + 1
    slot = torch.ops.cbx.slot_mapping(bt, rows, pos)
    assert torch.equal(slot, ops.slot_mapping(bt, rows, pos))
    kv_a, kv_b = (torch.zeros(n_blk * 65536, dtype=torch.uint8).cuda() for _ in range(2))
    qkv_a = torch.ops.cbx.layer_qkv(res, ln, wqkv, cs, pos, slot, kv_a)
    qkv_b = ops.layer_qkv(res, ln, wqkv, cs, pos, slot, kv_b)
    assert torch.equal(qkv_a[:, :1024], qkv_b[:, :1024]) and torch.equal(kv_a, kv_b) and kv_a.any()
    out_a = torch.ops.cbx.paged_attn_decode(qkv_a, kv_a, bt, rows, kv_len, 6)
    assert torch.equal(out_a, ops.paged_attn_decode(qkv_b, kv_b, bt, rows, kv_len, 6))
    r_a, r_b = res.clone(), res.clone()
    assert torch.ops.cbx.linear_residual(out_a, wo, r_a) is None
    ops.linear_residual(out_a, wo, r_b)
    assert torch.equal(r_a, r_b) and not torch.equal(r_a, res)
    assert torch.equal(torch.ops.cbx.rmsnorm(res, ln), ops.rmsnorm(res, ln))
    # the fused block (T <= 4) through the dispatcher
    r4, kv4 = res[:2].clone(), torch.zeros(n_blk * 65536, dtype=torch.uint8).cuda()
    r4b, kv4b = res[:2].clone(), kv4.clone()
    torch.ops.cbx.layer_attn_block(r4, ln, wqkv, wo, cs, pos[:2], slot[:2], kv4, bt, rows[:2], kv_len[:2], 6)
    ops.layer_attn_block(r4b, ln, wqkv, wo, cs, pos[:2], slot[:2], kv4b, bt, rows[:2], kv_len[:2], 6)
    assert torch.equal(r4, r4b) and torch.equal(kv4, kv4b) and not torch.equal(r4, res[:2])
    # ... and the fused MLP block
    g = torch.Generator().manual_seed(9)
    wgu = ops.pack_gate_up((torch.randn(8192, 1024, generator=g) * 0.03).to(BF16).cuda())
    wd_t = (torch.randn(4096, 1024, generator=g) * 0.02).to(BF16).cuda()
    m_a, m_b = res[:2].clone(), res[:2].clone()
    assert torch.ops.cbx.layer_mlp_block(m_a, ln, wgu, wd_t) is None
    ops.layer_mlp_block(m_b, ln, wgu, wd_t)
    assert torch.equal(m_a, m_b) and not torch.equal(m_a, res[:2])
