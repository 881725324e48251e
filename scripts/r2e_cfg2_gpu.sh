#!/bin/bash
# cfg2 (FastSpeech2 + ReferenceEncoder, no vocoder): bench line, conv breakdown, launch list
O=gpurun_out
timeout 600 python bench.py --workload cfg2 --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > $O/r2e_cfg2.json 2> $O/r2e_cfg2_conv_breakdown.txt
timeout 600 python bench.py --workload cfg2 --steps 1 --no-extra --no-cpu-baseline > $O/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/r2e_launches_cfg2_fp32.csv python bench.py --workload cfg2 --steps 1 --no-extra --no-cpu-baseline > $O/ncu1.log 2>&1
gzip -f $O/r2e_launches_cfg2_fp32.csv
cut -c1-300 $O/r2e_cfg2.json
