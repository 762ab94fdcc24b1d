#!/bin/bash
# Round-2 closing measurement pass (one box session): the bench line, the ncu evidence at the bench's own context, the
# greedy horizon with its near-tie gaps, and the reference arm's quick check.
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r02_bench_n1.err
timeout 1300 tools/gpu_profile_ctx440.sh
timeout 900 python tools/greedy_horizon.py > gpurun_out/r02_greedy_horizon.json 2> gpurun_out/r02_greedy_horizon.err; echo "horizon rc=$?"; head -c 1500 gpurun_out/r02_greedy_horizon.json
