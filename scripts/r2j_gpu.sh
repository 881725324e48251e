#!/bin/bash
O=gpurun_out
timeout 200 python scripts/resunit_one.py > $O/r2z_resunit_timings.txt 2>&1; cat $O/r2z_resunit_timings.txt
timeout 200 python scripts/resunit_one.py --ncu > $O/plain3.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:resunit -f -o /tmp/resunit python scripts/resunit_one.py --ncu > $O/ncu3.log 2>&1
ncu -i /tmp/resunit.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/r2z_resunit_raw.csv.gz
python scripts/ncu_line_stalls.py /tmp/resunit.ncu-rep resunit_bf16x3.cu 100 300 > $O/r2z_resunit_lines.txt 2>&1
