#!/bin/bash
# A/B of b = 1 knobs in one box session: prints p50 step ms per setting (8 timed M1 steps only; the b1 leg is what matters)
run() { env "$@" CBX_BENCH_ALLOW_KNOBS=1 python bench.py --no-cpu-baseline --no-e2e --no-fp8 --no-standin --steps 8 2>/dev/null | python -c "import json,sys;d=json.load(sys.stdin);print('$*', 'p50_ms', round(d['b1']['p50_ms'],4))"; }
for args in "$@"; do run $args; done
