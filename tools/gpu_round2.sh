#!/bin/bash
# Round-2 measurement pass (one box session): full bench line, switch-point sweep both ways, greedy horizon.
mkdir -p gpurun_out
python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_n1.err
python tools/bench_bsweep.py > gpurun_out/r02_bsweep_default.json 2>/dev/null; echo "sweep rc=$?"
CBX_GEMV_MAX_T=0 python tools/bench_bsweep.py > gpurun_out/r02_bsweep_tcgen05.json 2>/dev/null; echo "sweep tc rc=$?"
cat gpurun_out/r02_bsweep_default.json gpurun_out/r02_bsweep_tcgen05.json
python tools/greedy_horizon.py > gpurun_out/r02_greedy_horizon.json 2> gpurun_out/r02_greedy_horizon.err; echo "horizon rc=$?"; cat gpurun_out/r02_greedy_horizon.json
