#!/bin/bash
# ncu --set full captures of the opt-in two-launch T <= 4 layer (CBX_MLP_FUSED=1): the fused MLP block and the attention
# block's FOLD form, at b = 1, after a plain run of the same command
mkdir -p gpurun_out
export CBX_BENCH_ALLOW_KNOBS=1 CBX_MLP_FUSED=1
CMD1="python bench.py --pairs 1 --steps 2 --warmup 3 --ctx0 300 --no-cpu-baseline --no-e2e --no-b1 --no-fp8 --no-standin --no-probe"
T="timeout 300"
$T $CMD1 > gpurun_out/plain_b1_mlp_block.log 2>&1 || exit 1
$T ncu --set full --clock-control none --import-source on -k regex:layer_mlp_block -s 100 -c 2 -f -o gpurun_out/prof_b1_mlp_block $CMD1 > gpurun_out/ncu_b1_mlp_block.log 2>&1; echo "mlp rc=$?"
$T ncu --set full --clock-control none --import-source on -k regex:layer_attn_block -s 100 -c 2 -f -o gpurun_out/prof_b1_attn_block_fold $CMD1 > gpurun_out/ncu_b1_attn_block_fold.log 2>&1; echo "block rc=$?"
$T ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_B1_two_launch.csv $CMD1 > gpurun_out/ncu_launches_b1_two.log 2>&1; echo "list rc=$?"
