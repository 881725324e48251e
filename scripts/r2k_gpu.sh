#!/bin/bash
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_x3.py -x -q > $O/r2k_tests.log 2>&1; tail -2 $O/r2k_tests.log
timeout 200 python scripts/resunit_one.py 2>&1 | tee $O/r2k_resunit.txt
timeout 300 python scripts/x3_time.py 2>&1 | grep "c2" | tee $O/r2k_x3_c2.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline 2>/dev/null | cut -c1-220
