#!/bin/bash
O=gpurun_out
for cfg in "32 3" "128 3"; do
  set -- $cfg
  timeout 120 python scripts/x3_one.py $1 $2 > $O/plain.log 2>&1 || exit 1
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf16x3 -s 2 -c 2 -f -o /tmp/x3_$1_$2 python scripts/x3_one.py $1 $2 > $O/ncu_x3.log 2>&1
  ncu -i /tmp/x3_$1_$2.ncu-rep --page raw --csv > $O/r2g_x3_$1_$2_raw.csv 2>/dev/null
  python scripts/ncu_line_stalls.py /tmp/x3_$1_$2.ncu-rep tc_common.cuh 300 480 > $O/r2g_x3_$1_$2_drain_lines.txt 2>&1
  python scripts/ncu_line_stalls.py /tmp/x3_$1_$2.ncu-rep conv1d_bf16x3.cu 100 300 > $O/r2g_x3_$1_$2_kernel_lines.txt 2>&1
done
ls -la $O | tail
