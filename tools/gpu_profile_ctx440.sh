#!/bin/bash
# ncu evidence at the bench's own context (440): prompts of 436 rows put the first timed step there after 3 warm-up
# steps, so ncu has ~5 k launches in front of it instead of 54 k (the first version of this pass skipped 53 752 launches
# under ncu and ran out of its time limit).  (1) launch list of the whole run, the last step summarised;
# (2) --set full capture of three decode-attention launches of the timed steps.
mkdir -p gpurun_out
FLAGS="--no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin --no-probe"
CMD="python bench.py --steps 2 --warmup 3 --prefix 436 --ctx0 440 $FLAGS"
T="timeout 420"
$T $CMD > gpurun_out/plain_ctx440.log 2>&1 &&
$T ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list B128 @ctx440 rc=$?"
# attention launches: 30 per decode step (prefill uses another kernel); skip the 3 warm-up steps and the first timed step's first layers
$T ncu --set full --clock-control none --import-source on -k regex:paged_attn_decode_bal -s 100 -c 3 -f -o gpurun_out/prof_paged_attn_decode_ctx440 $CMD > gpurun_out/ncu_paged_attn_decode_ctx440.log 2>&1
echo "full paged_attn_decode_ctx440 rc=$?"
cp gpurun_out/plain_ctx440.log gpurun_out/plain_paged_attn_decode_ctx440.log
