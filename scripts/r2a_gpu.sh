#!/bin/bash
# round-2 first measurement pass: cfg3 per-shape conv breakdown in both precision modes, HiFi-GAN shape timings,
# launch list of a cfg3 step, ncu --set full of the HiFi-GAN conv shapes (raw page exported on the box: the report
# itself is too large to bring back).
set -x
O=gpurun_out
python bench.py --workload cfg3 --precision fp32 --conv-breakdown --no-extra --no-cpu-baseline --steps 5 > $O/r2a_cfg3_fp32.json 2> $O/r2a_cfg3_fp32_breakdown.txt
python bench.py --workload cfg3 --precision bf16 --conv-breakdown --no-extra --no-cpu-baseline --steps 5 > $O/r2a_cfg3_bf16.json 2> $O/r2a_cfg3_bf16_breakdown.txt
python scripts/hifigan_ncu.py both --time > $O/r2a_hifigan_shapes.txt 2>&1
python bench.py --workload cfg3 --precision fp32 --no-extra --no-cpu-baseline --steps 1 > $O/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2a_launches_cfg3_fp32.csv python bench.py --workload cfg3 --precision fp32 --no-extra --no-cpu-baseline --steps 1 > $O/ncu1.log 2>&1
gzip -f $O/r2a_launches_cfg3_fp32.csv
python scripts/hifigan_ncu.py both --few > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none -k regex:conv1d -o /tmp/r2a_hifigan python scripts/hifigan_ncu.py both --few > $O/ncu2.log 2>&1
ncu -i /tmp/r2a_hifigan.ncu-rep --page raw --csv > $O/r2a_hifigan_raw.csv 2>/dev/null
gzip -f $O/r2a_hifigan_raw.csv
ls -la $O /tmp/r2a_hifigan.ncu-rep
