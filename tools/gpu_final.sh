python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -c 200 gpurun_out/bench_full.err
CMD="python bench.py --steps 2 --warmup 3 --ctx0 204 --no-cpu-baseline --no-e2e --no-b1 --no-fp8"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:paged_attn_decode -s 160 -c 3 -f -o gpurun_out/prof_paged_attn_decode $CMD > gpurun_out/ncu_paged_attn_decode.log 2>&1; echo "attn capture rc=$?"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
