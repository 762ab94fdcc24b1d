"""Timeline inside the fused attention block (k_block.cu) at b = 1 (needs a -DCBX_BLOCK_PROF build:
make -C chatterbox_vllm_b200/csrc clean all EXTRA=-DCBX_BLOCK_PROF).  Runs the engine WITHOUT CUDA graphs so every
launch has its own stamp slot; prints, per layer of the last step, phase times (ns) relative to the first CTA's entry:
median / max over the 96 CTAs."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from chatterbox_vllm_b200 import _lib
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

L = int(os.environ.get("LAYERS", "30"))
dev = torch.device("cuda:0")
eng = T3Engine(random_state_dict(L, device=dev, seed=0), device=dev, max_pairs=1, max_model_len=700, use_graph=False)
sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=400, ignore_eos=True)
reqs = [Request(random_prompt(40, seed=1), random_cond_emb(2, pinned=True), sp, uid=7)]
lib = _lib.load()
lib.cbx_block_prof_read.restype = ctypes.c_int
buf = np.zeros((128, 128, 12), dtype=np.uint64)
names = ["entry", "pre-wait issued", "past wait", "rows normed", "QKV done", "exchange bar", "attn loop", "partial bar",
         "merged", "O+ticket", "end"]
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    eng._init_sampler_state(reqs, P)
    eng.prefill(reqs, P)
    eng.decode(1, 250, 400)
    torch.cuda.synchronize()
    lib.cbx_block_prof_read(buf.ctypes.data_as(ctypes.c_void_p), 128)
    eng.decode(1, 2, 400)
    torch.cuda.synchronize()
    n = lib.cbx_block_prof_read(buf.ctypes.data_as(ctypes.c_void_p), 128)
for cl in (4, 6, 8):
    lib.cbx_debug_block_max_clusters.restype = ctypes.c_int
    print(f"max co-resident clusters of {cl} CTAs (TT=2 config):", lib.cbx_debug_block_max_clusters(2, cl))
st = buf[L + 3, :96].astype(np.int64)
late = (st[:, 0] - st[:, 0].min()) > 4000
print("layer 3: CTAs entering > 4 us after the first:", int(late.sum()), "clusters (heads) affected:", sorted(set(np.nonzero(late)[0] // 6)))
print(f"{n} block launches recorded (2 steps x {L} layers)")
prev_end = None
for i in range(L, min(n, 2 * L)):
    st = buf[i, :96].astype(np.int64)
    t0 = st[:, 0].min()
    gap = "" if prev_end is None else f" gap since previous block's end {t0 - prev_end:6d} ns"
    row = " | ".join(f"{names[k]} {int(np.median(st[:, k] - t0)):5d}/{int((st[:, k] - t0).max()):5d}" for k in range(11))
    if i in (L, L + 1, L + 2, L + L // 2, 2 * L - 1):
        print(f"layer {i - L:2d}:{gap}\n    {row}")
    prev_end = st[:, 10].max()
dur = [int(buf[i, :96].astype(np.int64)[:, 10].max() - buf[i, :96].astype(np.int64)[:, 2].min()) for i in range(L, min(n, 2 * L))]
span = [int(buf[i + 1, :96].astype(np.int64)[:, 0].min() - buf[i, :96].astype(np.int64)[:, 0].min()) for i in range(L, min(n, 2 * L) - 1)]
print("block past-wait -> end, ns: median", int(np.median(dur)), " layer period (entry to entry), ns: median", int(np.median(span)))
