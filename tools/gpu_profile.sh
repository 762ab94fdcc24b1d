#!/bin/bash
# ncu evidence for the bench step: (1) launch lists with device times (B = 128 pairs and B = 1), (2) --set full
# captures of the dominant kernels.  Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1 --no-fp8"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
CMD1="python bench.py --pairs 1 --steps 2 --warmup 3 --ctx0 300 --no-cpu-baseline --no-e2e --no-b1 --no-fp8"
$CMD1 > gpurun_out/plain_b1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_B1.csv $CMD1 > gpurun_out/ncu_launches_b1.log 2>&1
echo "launch list B1 rc=$?"
prof() { # name, kernel regex, launches to skip, launches to capture
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c $4 -f -o gpurun_out/prof_$1 $CMD > gpurun_out/ncu_$1.log 2>&1
  echo "full $1 rc=$?"
}
prof paged_attn_decode paged_attn_decode 160 3
prof gemm_tc gemm_tc 1000 4                 # decode-sized (M = 256): QKV / O / gate_up / down of one layer
prof gemm_tc_prefill gemm_tc 300 4          # prefill-sized (M = 16000): tensor-pipe utilisation
prof paged_attn_prefill paged_attn_prefill 40 2
prof cfg_sample cfg_sample 2 3
# FP8-e4m3 weight path (reported separately): launch list of the same step on FP8 weights + the W8 GEMM form
CMD8="python bench.py --weights fp8_e4m3 --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1"
$CMD8 > gpurun_out/plain_fp8.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_fp8.csv $CMD8 > gpurun_out/ncu_launches_fp8.log 2>&1
echo "launch list fp8 rc=$?"
$CMD8 > gpurun_out/plain_fp8.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 1000 -c 4 -f -o gpurun_out/prof_gemm_tc_fp8w $CMD8 > gpurun_out/ncu_gemm_tc_fp8w.log 2>&1
echo "full gemm_tc_fp8w rc=$?"
