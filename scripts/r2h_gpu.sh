#!/bin/bash
# full GPU suite + headline and cfg2 bench lines
O=gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > $O/r2h_tests.log 2>&1; echo "tests rc=$?" >> $O/r2h_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline --conv-breakdown > $O/r2h_bench.log 2> $O/r2h_breakdown.txt
timeout 600 python bench.py --workload cfg2 --steps 10 --warmup 3 --no-extra --no-cpu-baseline --conv-breakdown > $O/r2h_bench_cfg2.log 2> $O/r2h_breakdown_cfg2.txt
tail -5 $O/r2h_tests.log; cut -c1-250 $O/r2h_bench.log | tail -1;  cut -c1-250 $O/r2h_bench_cfg2.log | tail -1
