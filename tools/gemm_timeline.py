"""Timeline of the tcgen05 GEMM launches of one decode step (needs a -DCBX_GEMM_PROF build:
make -C chatterbox_vllm_b200/csrc clean all EXTRA=-DCBX_GEMM_PROF).  Runs the engine WITHOUT CUDA graphs for a
few steps so that every launch gets its own stamp slot; prints, per launch of the last step, the stamps
relative to the previous GEMM's last epilogue end (ns): first entry, last past-wait, first-stage landed,
last MMA issued, accumulator complete, last epilogue done."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from chatterbox_vllm_b200 import _lib
from chatterbox_vllm_b200.engine import Request, SamplingParams, T3Engine
from chatterbox_vllm_b200.synthetic import random_cond_emb, random_prompt, random_state_dict

B, L = int(os.environ.get("PAIRS", "128")), int(os.environ.get("LAYERS", "4"))
dev = torch.device("cuda:0")
eng = T3Engine(random_state_dict(L, device=dev, seed=0), device=dev, max_pairs=B, max_model_len=400, max_rows=16384,
               use_graph=False)
sp = SamplingParams(temperature=0.5, top_p=0.5, repetition_penalty=2.0, max_tokens=64, ignore_eos=True)
reqs = [Request(random_prompt(165, seed=i), random_cond_emb(i, pinned=True), sp, uid=1 + i) for i in range(B)]
lib = _lib.load()
lib.cbx_gemm_prof_read.restype = ctypes.c_int
buf = np.zeros((256, 160, 8), dtype=np.uint64)
with torch.cuda.stream(eng.stream):
    P, held = eng._admit(reqs)
    eng._init_sampler_state(reqs, P)
    eng.prefill(reqs, P)
    eng.decode(B, 3, 256)
    torch.cuda.synchronize()
    lib.cbx_gemm_prof_read(buf.ctypes.data_as(ctypes.c_void_p), 256)     # restart ids
    sys.stderr.write("---- measured step\n")
    eng.decode(B, 1, 256)
    torch.cuda.synchronize()
    n = lib.cbx_gemm_prof_read(buf.ctypes.data_as(ctypes.c_void_p), 256)
print(f"{n} GEMM launches in the step")
prev_end = None
for i in range(min(n, 256)):
    st = buf[i].astype(np.int64)
    live = st[:, 0] > 0
    if not live.any():
        continue
    s = st[live]
    t0 = s[:, 0].min()
    base = prev_end if prev_end is not None else t0
    rel = lambda v: int(v - base)
    print(f"launch {i:3d} ctas={live.sum():3d}: entry {rel(t0):7d} | past wait {rel(s[:, 2].max()):7d} | 1st stage {rel(s[s[:, 3] > 0, 3].min()):7d}.."
          f"{rel(s[s[:, 3] > 0, 3].max()):7d} | MMA issued {rel(s[s[:, 4] > 0, 4].max()):7d} | acc done {rel(s[:, 5].max()):7d} | "
          f"epi done {rel(s[:, 6].min()):7d}..{rel(s[:, 6].max()):7d}   dur {int(s[:, 6].max() - s[:, 2].min())} ns")
    prev_end = s[:, 6].max()
