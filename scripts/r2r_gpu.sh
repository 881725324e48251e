#!/bin/bash
O=gpurun_out
python scripts/panel_ncu.py 32 11 > $O/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:panel -s 1 -c 1 -o /tmp/p32 python scripts/panel_ncu.py 32 11 > $O/ncu3.log 2>&1
python scripts/ncu_line_stalls.py /tmp/p32.ncu-rep conv1d_tf32_panel.cu 150 360 --cols > $O/r2r_stalls_panel_32_11.txt 2>&1
