"""Decode attention: uniform vs ragged context lengths at the same total KV bytes.  python tools/bench_attn_ragged.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_vllm_b200 import ops

dev = torch.device("cuda:0")
T, L = 256, 30
g = torch.Generator().manual_seed(0)
for name, lens in (("uniform 450", torch.full((T,), 450)),
                   ("ragged U[200,700]", torch.randint(200, 701, (T,), generator=g)),
                   ("ragged U[50,850]", torch.randint(50, 851, (T,), generator=g)),
                   ("bimodal 100/800", torch.where(torch.arange(T) % 2 == 0, torch.tensor(100), torch.tensor(800)))):
    lens = lens.to(torch.int32)
    n_pages = (lens + 15) // 16
    maxp = int(n_pages.max())
    nb = int(n_pages.sum()) + 1
    bt = torch.zeros(T, maxp, dtype=torch.int32)
    nxt = 1
    for r in range(T):
        bt[r, :n_pages[r]] = torch.arange(nxt, nxt + int(n_pages[r]), dtype=torch.int32)
        nxt += int(n_pages[r])
    n_kv = 4
    kv = torch.zeros(n_kv * nb * 65536, dtype=torch.uint8, device=dev)
    qkv = (torch.randn(T, 3072) * 1.0).to(torch.bfloat16).to(dev)
    out = torch.empty(T, 1024, dtype=torch.bfloat16, device=dev)
    rows = torch.arange(T, dtype=torch.int32, device=dev)
    bt_d, kvl = bt.to(dev), lens.to(dev)
    mx = int(lens.max())
    fn = lambda i: ops.paged_attn_decode(qkv, kv[(i % n_kv) * nb * 65536:(i % n_kv + 1) * nb * 65536], bt_d, rows, kvl, mx, out=out)  # noqa: E731
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for i in range(L):
            fn(i)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for i in range(L):
                fn(i)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            gr.replay()
        e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / (5 * L) * 1e3
    byt = float(lens.sum()) * 4096 + T * 4096
    print(f"{name:22s}: {us:7.1f} us/launch (attention + merge), {byt / us / 1e3:6.0f} GB/s, total pages {int(n_pages.sum())}", flush=True)
