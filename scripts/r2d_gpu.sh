#!/bin/bash
# 8-epilogue-warp CTA-pair kernel: parity tests, per-shape timing, headline bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_tc.py -x -q -m gpu > gpurun_out/r2d_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2d_tests.log
timeout 300 python scripts/hifigan_ncu.py fp32 --time > gpurun_out/r2d_shapes.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/r2d_bench.log 2>&1
tail -3 gpurun_out/r2d_tests.log; cat gpurun_out/r2d_shapes.log | head -30; tail -1 gpurun_out/r2d_bench.log | cut -c1-400
