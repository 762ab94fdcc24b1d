#!/bin/bash
# ncu evidence for the bench step: (1) launch list with device times, (2) --set full captures of
# the dominant kernels.  Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
prof() { # kernel regex, launches to skip
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 3 -f -o gpurun_out/prof_$1 $CMD > gpurun_out/ncu_$1.log 2>&1
  echo "full $1 rc=$?"
}
prof paged_attn_decode 160
prof gemm_tc 1000
prof cfg_sample 2
