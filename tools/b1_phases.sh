#!/bin/bash
# In-graph marginal cost of the fused attention block's phases at b = 1: p50 step with every CTA of the block returning
# after phase k (CBX_BLOCK_STOP=k; results are garbage), and of the two GEMV kernels that follow it (CBX_SKIP bits 8 / 16
# do not apply on the fused path, so those come from the unfused run's numbers in profiles/).
mkdir -p gpurun_out
for k in 1 2 3 4 5 6 0; do
  CBX_BLOCK_STOP=$k CBX_BENCH_ALLOW_KNOBS=1 python bench.py --no-cpu-baseline --no-e2e --no-fp8 --steps 8 2>/dev/null \
    | python -c "import json,sys;d=json.load(sys.stdin);print('stop=$k p50_ms', round(d['b1']['p50_ms'],4))"
done
CBX_BLOCK_FUSED=0 CBX_BENCH_ALLOW_KNOBS=1 python bench.py --no-cpu-baseline --no-e2e --no-fp8 --steps 8 2>/dev/null | python -c "import json,sys;d=json.load(sys.stdin);print('unfused p50_ms', round(d['b1']['p50_ms'],4))"
