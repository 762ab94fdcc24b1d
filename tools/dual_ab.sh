#!/bin/bash
# A/B of the two-stream decode step in one box session: M1 ms/step per setting (short runs: the device-timed value only)
run() { env "$@" CBX_BENCH_ALLOW_KNOBS=1 python bench.py --no-cpu-baseline --no-e2e --no-fp8 --no-standin --no-b1 --no-probe --steps 60 2>/dev/null | python -c "import json,sys;d=json.load(sys.stdin);print('$*', 'ms_per_step', round(d['ms_per_step'],4), 'tok/s', round(d['value']), 'launches', d.get('gpu_launches'))"; }
for args in "$@"; do run $args; done
