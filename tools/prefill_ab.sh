#!/bin/bash
# prefill leg of the bench at different step-buffer capacities (chunk sizes): ms and fraction of the bf16 peak
for mr in "$@"; do
  python bench.py --no-cpu-baseline --no-e2e --no-fp8 --no-standin --no-b1 --no-probe --steps 8 --max-rows $mr 2>/dev/null | python -c "import json,sys;d=json.load(sys.stdin);p=d['prefill'];print('max_rows $mr prefill_ms', round(p['ms'],2), 'frac', round(p['frac_of_bf16_peak'],3), 'step_ms', round(d['ms_per_step'],4))"
done
