#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r2d_pytest.txt 2>&1
tail -15 $O/r2d_pytest.txt
python bench.py --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > $O/r2d_bench_cfg3_fp32.json 2> $O/r2d_cfg3_fp32_breakdown.txt
python bench.py --steps 10 --precision bf16 --conv-breakdown --no-extra --no-cpu-baseline > $O/r2d_bench_cfg3_bf16.json 2> $O/r2d_cfg3_bf16_breakdown.txt
python scripts/hifigan_ncu.py fp32 --time > $O/r2d_hifigan_shapes.txt 2>&1
