#!/bin/bash
O=gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > $O/r2i_tests.log 2>&1; echo "tests rc=$?" >> $O/r2i_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --precision bf16 --no-extra --no-cpu-baseline --conv-breakdown > $O/r2i_bench_bf16.log 2> $O/r2i_breakdown_bf16.txt
tail -5 $O/r2i_tests.log; cut -c1-250 $O/r2i_bench_bf16.log | tail -1; grep "^step" $O/r2i_breakdown_bf16.txt; grep "bf16 mode" $O/parity_report.txt
