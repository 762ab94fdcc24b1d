"""Repeat small balanced-attention launches and compare bits with the first result (race hunting)."""
import sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from chatterbox_vllm_b200 import ops
from test_gpu_kernels import _build_paged, _ref_attn, dev
from oracle.t3_ref import T3ModelConfig
cfg = T3ModelConfig()
for lens in ([70, 75, 81, 90, 96, 100], [70 + (i * 5) % 40 for i in range(64)]):
    kv, bt_d, qkv_all, cs = _build_paged(ops, cfg, lens, seed=3)
    S = len(lens)
    q = torch.randn(S, 3072, generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    args = (dev(q), kv, bt_d, dev(torch.arange(S, dtype=torch.int32)), dev(torch.tensor(lens, dtype=torch.int32)), max(lens))
    first = ops.paged_attn_decode(*args).clone()
    bad = 0
    for i in range(500):
        o = ops.paged_attn_decode(*args)
        if not torch.equal(o, first):
            bad += 1
    err = max((first[r].float().cpu() - _ref_attn(cfg, qkv_all[r], q[r, :1024].float(), lens[r], cs)).abs().max().item() for r in range(S))
    print(f"T={S}: {bad}/500 launches differ from the first; max |err| vs oracle {err:.4f}")
