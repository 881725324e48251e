#!/bin/bash
O=gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -x -q -k "panel or ragged or hifigan or vocoder or cfg3" > $O/r2u_pytest.txt 2>&1; tail -3 $O/r2u_pytest.txt
: > $O/r2u_panel_time.txt
for shape in "32 11" "32 7" "32 3" "32 11 5"; do python scripts/panel_time.py $shape >> $O/r2u_panel_time.txt 2>&1; done
cat $O/r2u_panel_time.txt
python bench.py --steps 10 --conv-breakdown --no-extra --no-cpu-baseline > $O/r2u_bench_cfg3_fp32.json 2> $O/r2u_cfg3_fp32_breakdown.txt
