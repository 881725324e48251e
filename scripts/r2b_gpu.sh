#!/bin/bash
set -x
O=gpurun_out
python scripts/desc_shift_probe.py > $O/r2b_desc_probe.txt 2>&1
python bench.py --steps 5 > $O/r2b_bench_default.json 2> $O/r2b_bench_default.err
for shape in "32 32 3" "32 32 11" "64 64 3" "128 128 3"; do
  TRACE_B=4 TRACE_S=40960 TRACE_ROWS=200 python scripts/conv_trace.py ts $shape > "$O/r2b_trace_ts_${shape// /_}.txt" 2>&1
done
