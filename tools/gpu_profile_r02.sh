#!/bin/bash
# Round-2 ncu evidence, each capture after a plain run of the same command that exited 0:
#  (1) launch lists (device time per launch) of one M1 decode step at the bench's own context (440) and of one b = 1 step
#      on the fused-block path and on the five-kernel chain it replaced;
#  (2) --set full captures of the dominant kernel at that context (decode attention), of the decode GEMMs, of every b = 1
#      kernel (the fused attention block, the three GEMV shapes that follow it, the five-kernel chain's QKV / attention / O
#      kernels), of the split-K add+RMSNorm kernel and of the sampler.
mkdir -p gpurun_out
FLAGS="--no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin"
CMD="python bench.py --steps 2 --warmup 3 --ctx0 440 $FLAGS"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list B128 rc=$?"
CMD1="python bench.py --pairs 1 --steps 2 --warmup 3 --ctx0 300 $FLAGS"
$CMD1 > gpurun_out/plain_b1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_B1.csv $CMD1 > gpurun_out/ncu_launches_b1.log 2>&1
echo "launch list B1 (fused block) rc=$?"
export CBX_BENCH_ALLOW_KNOBS=1
CBX_BLOCK_FUSED=0 $CMD1 > gpurun_out/plain_b1c.log 2>&1 &&
CBX_BLOCK_FUSED=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_B1_chain.csv $CMD1 > gpurun_out/ncu_launches_b1c.log 2>&1
echo "launch list B1 (five-kernel chain) rc=$?"
unset CBX_BENCH_ALLOW_KNOBS
prof() { # name, kernel regex, launches to skip, launches to capture, command...
  local name=$1 re=$2 skip=$3 cnt=$4; shift 4
  "$@" > gpurun_out/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$re -s $skip -c $cnt -f -o gpurun_out/prof_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "full $name rc=$?"
}
prof paged_attn_decode_ctx440 paged_attn_decode 160 3 $CMD
prof gemm_tc_decode gemm_tc 1000 4 $CMD
prof rmsnorm_reduce4 rmsnorm_reduce4 200 2 $CMD
prof cfg_sample cfg_sample 2 2 $CMD
prof b1_attn_block layer_attn_block 100 2 $CMD1
prof b1_gemv gemv_fused 300 3 $CMD1                  # gate_up (accumulator prologue), down (+ accumulator clear), next layer ... head
export CBX_BENCH_ALLOW_KNOBS=1 CBX_BLOCK_FUSED=0
prof b1_chain_gemv gemv_fused 400 4 $CMD1            # QKV, O, gate_up, down of the five-kernel chain
prof b1_chain_attn_small paged_attn_small 100 2 $CMD1
