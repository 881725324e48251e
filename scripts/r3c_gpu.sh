#!/bin/bash
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r3c_pytest.txt 2>&1; tail -3 $O/r3c_pytest.txt
python bench.py --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > $O/r3c_bench_cfg3_fp32.json 2> $O/r3c_cfg3_fp32_breakdown.txt
