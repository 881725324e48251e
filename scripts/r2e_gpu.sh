#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests/test_gpu_sharded.py -x -q > $O/r2e_pytest.txt 2>&1; tail -3 $O/r2e_pytest.txt
python scripts/panel_ncu.py 32 11 > $O/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:panel -s 1 -c 1 -o /tmp/p32 python scripts/panel_ncu.py 32 11 > $O/ncu3.log 2>&1
python scripts/ncu_lines.py /tmp/p32.ncu-rep 60 > $O/r2e_lines_panel_32_11.txt 2>&1
ncu -i /tmp/p32.ncu-rep --page raw --csv > $O/r2e_raw_panel_32_11.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:panel -s 1 -c 1 -o /tmp/p64 python scripts/panel_ncu.py 64 11 > $O/ncu4.log 2>&1
python scripts/ncu_lines.py /tmp/p64.ncu-rep 60 > $O/r2e_lines_panel_64_11.txt 2>&1
ncu -i /tmp/p64.ncu-rep --page raw --csv > $O/r2e_raw_panel_64_11.csv 2>/dev/null
ls -la /tmp/*.ncu-rep
