#!/bin/bash
# ncu evidence for the bench step: (1) launch list with device times, (2) --set full captures of
# the dominant kernels.  Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
for k in ${KERNELS:-paged_attn_decode gemm_tc}; do
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s ${SKIP:-600} -c 3 -f -o gpurun_out/prof_$k $CMD > gpurun_out/ncu_$k.log 2>&1
  echo "full $k rc=$?"
done
