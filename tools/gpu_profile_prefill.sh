#!/bin/bash
# ncu launch list of the M1 prefill (prefix 200, 4 chunks of 12 800 rows, 30 layers): the first ~2700 launches of a bench run
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin --no-probe"
timeout 300 $CMD > gpurun_out/plain_prefill.log 2>&1 || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 2700 --csv --log-file gpurun_out/launches_prefill.csv $CMD > gpurun_out/ncu_launches_prefill.log 2>&1
echo "prefill launch list rc=$?"
