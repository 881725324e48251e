#!/bin/bash
# round-2 final measurement pass (one GPU): default bench line, bf16 line, launch lists, ncu --set full of the HiFi-GAN
# conv shapes (raw page exported on the box), shape timings
set -x
O=gpurun_out
python bench.py > $O/r2final_bench_default.json 2> $O/r2final_bench_default.err
python bench.py --precision bf16 --no-extra --no-cpu-baseline > $O/r2final_bench_bf16.json 2>> $O/r2final_bench_default.err
python bench.py --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > /dev/null 2> $O/r2final_cfg3_fp32_conv_breakdown.txt
python bench.py --steps 10 --precision bf16 --conv-breakdown --no-extra --no-cpu-baseline > /dev/null 2> $O/r2final_cfg3_bf16_conv_breakdown.txt
python scripts/hifigan_ncu.py both --time > $O/r2final_hifigan_shape_timings.txt 2>&1
python bench.py --steps 1 --no-extra --no-cpu-baseline > $O/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/r2final_launches_cfg3_fp32.csv python bench.py --steps 1 --no-extra --no-cpu-baseline > $O/ncu1.log 2>&1
gzip -f $O/r2final_launches_cfg3_fp32.csv
python scripts/hifigan_ncu.py fp32 --few > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none -k regex:conv1d -o /tmp/r2final_hifigan python scripts/hifigan_ncu.py fp32 --few > $O/ncu2.log 2>&1
ncu -i /tmp/r2final_hifigan.ncu-rep --page raw --csv > $O/r2final_hifigan_raw.csv 2>/dev/null
gzip -f $O/r2final_hifigan_raw.csv
ls -la $O | tail -12
