#!/bin/bash
# Round-2 ncu evidence, each capture after a plain run of the same command that exited 0 (every ncu run under its own
# timeout).  ncu counts kernel nodes of graph launches, so a decode step at the bench's own context is reached by
# SKIPPING the launches before it: 2606 launches of set-up + two prefills, then 214 per M1 step, 239 positioning steps.
#  (1) launch list (device time per launch) of two M1 decode steps at context 440;
#  (2) --set full captures of the dominant kernel at that context (decode attention), of the decode GEMMs, of the split-K
#      add+RMSNorm kernel, and of every b = 1 kernel: the fused attention block, the GEMV shapes that follow it, and the
#      five-kernel chain's QKV / O GEMVs and small attention kernel (CBX_BLOCK_FUSED=0);
#  (3) launch list of a b = 1 step on the five-kernel chain (the fused path's list: launches_B1.csv of the same script's
#      first version, kept).
mkdir -p gpurun_out
FLAGS="--no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin"
CMD="python bench.py --steps 2 --warmup 3 --ctx0 440 $FLAGS"
CMD1="python bench.py --pairs 1 --steps 2 --warmup 3 --ctx0 300 $FLAGS"
T="timeout 300"
$T $CMD > gpurun_out/plain.log 2>&1 &&
$T ncu --metrics gpu__time_duration.sum --clock-control none -s 53752 -c 430 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list B128 @ctx440 rc=$?"
prof() { # name, kernel regex, launches to skip, launches to capture, command...
  local name=$1 re=$2 skip=$3 cnt=$4; shift 4
  $T "$@" > gpurun_out/plain_$name.log 2>&1 &&
  $T ncu --set full --clock-control none --import-source on -k regex:$re -s $skip -c $cnt -f -o gpurun_out/prof_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "full $name rc=$?"
}
prof paged_attn_decode_ctx440 paged_attn_decode_bal 7170 3 $CMD
prof gemm_tc_decode gemm_tc 1000 4 $CMD
prof rmsnorm_reduce4 rmsnorm_reduce4 200 2 $CMD
prof b1_attn_block layer_attn_block 100 2 $CMD1
prof b1_gemv gemv_fused 300 3 $CMD1                  # gate_up (accumulator prologue), down (+ accumulator clear), gate_up of the next layer
export CBX_BENCH_ALLOW_KNOBS=1 CBX_BLOCK_FUSED=0
prof b1_chain_gemv gemv_fused 400 4 $CMD1            # QKV, O, gate_up, down of the five-kernel chain
prof b1_chain_attn_small paged_attn_small 100 2 $CMD1
$T $CMD1 > gpurun_out/plain_b1c.log 2>&1 &&
$T ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_B1_chain.csv $CMD1 > gpurun_out/ncu_launches_b1c.log 2>&1
echo "launch list B1 (five-kernel chain) rc=$?"
