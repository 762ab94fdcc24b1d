#!/bin/bash
# ncu --set full capture (with source) of the prefill QKV+RoPE+append GEMM (CTA-pair form, 256-wide tiles) inside an M1 prefill
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin --no-probe"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 400 -c 8 -f -o gpurun_out/prof_prefill_gemms $CMD > gpurun_out/ncu_prefill_gemms.log 2>&1
echo "prefill gemms rc=$?"
