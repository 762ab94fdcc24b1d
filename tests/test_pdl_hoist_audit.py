"""Static check on the built library (no GPU): no kernel may read what its predecessor produces
before griddepcontrol.wait.  nvcc is free to move `const __restrict__` loads above the wait's memory
clobber (it did, for the attention page-range plan: LDG.E.CONSTANT before ACQBULK, stale plans under
programmatic dependent launch), so the SASS is audited: only the kernels that prefetch WEIGHTS ahead
of the wait by design may have global loads in front of it."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "chatterbox_vllm_b200", "libcbx_t3.so")
# weight prefetch before the wait is the point of these (k_gemv.cu / k_gemm.cu headers); the stream-ahead GEMV
# experiment issues its weight rows as bulk copies (UBLKCP) there
WEIGHT_PREFETCH = ("gemv_fused_kernel", "gemv_stream_kernel", "gemm_tc_kernel")
# the b = 1 attention kernel reads its page ids (block table via row_seq: written at admission / re-staged with the same
# value every step, never by the kernels it waits for) ahead of the wait: exactly two loads, checked below
PAGE_ID_PREFETCH = "paged_attn_small_kernel"
# the 4-slice add+RMSNorm kernel reads its norm weights (one 8-byte load) ahead of the wait
NORM_WEIGHT_PREFETCH = "rmsnorm_reduce4_kernel"
# the fused attention block (k_block.cu) requests its Wqkv rows (8 loads), Wo slice (3), norm weights (4) and its rows'
# page ids (row_seq -> block table: 2) ahead of the wait: exactly 17 loads
BLOCK_PREFETCH = "layer_attn_block_kernel"
# its FOLD form (after the fused MLP block) fetches the norm weights in the chunk mapping: 1 load instead of 4 = 14
# the fused MLP block (k_mlp.cu): gate / up rows of the warp's two features (16 coherent loads, ordered against the wait),
# norm weights (1) and the bulk copy of its WdT rows = 18
MLP_PREFETCH = "layer_mlp_block_kernel"


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(SO), reason="needs cuobjdump and the built library")
def test_no_activation_load_hoisted_above_griddepcontrol_wait():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "audit_pdl_hoist.py"), SO],
                         capture_output=True, text=True, check=True).stdout
    flagged = [l for l in out.splitlines() if "before the wait" in l and not l.startswith(WEIGHT_PREFETCH)
               and not (l.startswith(PAGE_ID_PREFETCH) and ": 2 load(s)" in l)
               and not (l.startswith(NORM_WEIGHT_PREFETCH) and ": 1 load(s)" in l)
               and not (l.startswith(BLOCK_PREFETCH) and ", false>: 17 load(s)" in l)
               and not (l.startswith(BLOCK_PREFETCH) and ", true>: 14 load(s)" in l)
               and not (l.startswith(MLP_PREFETCH) and ": 18 load(s)" in l)]
    assert not flagged, "loads hoisted above griddepcontrol.wait:\n" + "\n".join(flagged)
    assert any(l.startswith("paged_attn_decode_bal_kernel") and l.endswith("clean") for l in out.splitlines()), out[:400]


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(SO), reason="needs cuobjdump and the built library")
def test_no_shared_memory_read_in_flight_when_a_ring_stage_is_released():
    """Every mbarrier arrive that hands a shared-memory stage back to the TMA producer must come after the warp's
    LDS / LDSM of that stage have been PERFORMED (their results consumed, or a MEMBAR in between).  ptxas once
    scheduled the prefill-attention release above the MMAs consuming the last ldmatrix; the refill then overtook
    the read and a few heads saw the next page (tools/audit_stage_release.py walks the SASS)."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "audit_stage_release.py"), SO], capture_output=True, text=True)
    assert r.returncode == 0 and "UNSAFE" not in r.stdout, r.stdout
    for k in ("paged_attn_prefill_kernel", "paged_attn_decode_bal_kernel", "paged_attn_decode_kernel<true>"):
        assert any(l.startswith(k) and ": ok" in l for l in r.stdout.splitlines()), r.stdout[:600]
