#!/bin/bash
# round-2 final measurement pass (one GPU): tests, smoke, default bench + reference arm, bf16 line, conv breakdowns, shape
# timings, launch list, ncu --set full of the split-bf16 kernel at the four HiFi-GAN stage shapes
O=gpurun_out
python -m pytest tests -m gpu -q > $O/r2z_pytest.txt 2>&1; tail -3 $O/r2z_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2z_smoke.txt 2>&1; tail -2 $O/r2z_smoke.txt
python bench.py > $O/r2z_bench_default.json 2> $O/r2z_bench_default.err; tail -2 $O/r2z_bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r2z_bench_reference.json 2>&1
python bench.py --precision bf16 --no-extra --no-cpu-baseline > $O/r2z_bench_bf16.json 2>> $O/r2z_bench_default.err
python bench.py --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > /dev/null 2> $O/r2z_cfg3_fp32_conv_breakdown.txt
python bench.py --steps 10 --precision bf16 --conv-breakdown --no-extra --no-cpu-baseline > /dev/null 2> $O/r2z_cfg3_bf16_conv_breakdown.txt
python bench.py --workload cfg2 --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > $O/r2z_bench_cfg2.json 2> $O/r2z_cfg2_fp32_conv_breakdown.txt
python scripts/x3_time.py > $O/r2z_x3_shape_timings.txt 2>&1
python bench.py --steps 1 --no-extra --no-cpu-baseline > $O/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/r2z_launches_cfg3_fp32.csv python bench.py --steps 1 --no-extra --no-cpu-baseline > $O/ncu1.log 2>&1
gzip -f $O/r2z_launches_cfg3_fp32.csv
for cfg in "32 3" "32 11" "64 11" "128 3" "128 11" "256 11"; do
  set -- $cfg
  timeout 120 python scripts/x3_one.py $1 $2 > $O/plain2.log 2>&1 || continue
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf16x3 -s 8 -c 4 -f -o /tmp/x3_$1_$2 python scripts/x3_one.py $1 $2 > $O/ncu2.log 2>&1
  ncu -i /tmp/x3_$1_$2.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/r2z_x3_$1_$2_raw.csv.gz
done
python scripts/resunit_one.py > $O/r2z_resunit_timings.txt 2>&1
timeout 200 python scripts/resunit_one.py --ncu > $O/plain3.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:resunit -f -o /tmp/resunit python scripts/resunit_one.py --ncu > $O/ncu3.log 2>&1
ncu -i /tmp/resunit.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/r2z_resunit_raw.csv.gz
ls -la $O | tail -30
