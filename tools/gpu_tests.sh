#!/bin/bash
# Runs the GPU parity tests group by group, each in its own process under a timeout, so a
# hung kernel in one group cannot take the rest (or the box) with it.  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, pytest args...
  local name=$1; shift
  timeout -k 5 ${TMO:-420} python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/test_$name.log 2>&1
  echo "$name rc=$? $(tail -1 gpurun_out/test_$name.log)"
}
run ints    tests/test_gpu_kernels.py -k "slot_mapping or rmsnorm"
run gemv    tests/test_gpu_kernels.py -k "gemv"
run gemm    tests/test_gpu_kernels.py -k "gemm_tcgen05"
run silu    tests/test_gpu_kernels.py -k "silu"
run rope    tests/test_gpu_kernels.py -k "rope and not fused"
run attn    tests/test_gpu_kernels.py -k "paged_attention"
run fused   tests/test_gpu_kernels.py -k "fused or residual or split_k"
run post    tests/test_gpu_kernels.py -k "postprocess or cond_encoder"
run sampler tests/test_gpu_engine.py -k "sampler or prefill_embed"
run engine  tests/test_gpu_engine.py -k "engine"
run facade  tests/test_gpu_facade.py
run fp8     tests/test_gpu_fp8.py
run block   tests/test_gpu_block.py tests/test_gpu_mlp_block.py
run sched   tests/test_gpu_scheduler.py
run configs tests/test_gpu_configs.py
run torchop tests/test_torch_ops.py
run mtl     tests/test_mtl_surface.py
run shard   tests/test_gpu_sharding_nccl.py
run noalloc tests/test_zz_no_alloc.py
