#!/bin/bash
# split-bf16 ResBlocks in the fp32 mode: parity tests and the headline bench
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_sharded.py tests/test_gpu_x3.py -q -m gpu > $O/r2f_tests.log 2>&1; echo "tests rc=$?" >> $O/r2f_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline --conv-breakdown > $O/r2f_bench.log 2> $O/r2f_breakdown.txt
tail -25 $O/r2f_tests.log; tail -1 $O/r2f_bench.log | cut -c1-300; cat $O/parity_report.txt | tail -30
