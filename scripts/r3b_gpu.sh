#!/bin/bash
O=gpurun_out
python -m pytest tests/test_gpu_kernels.py -x -q -k "panel or ragged" > $O/r3b_pytest.txt 2>&1; tail -3 $O/r3b_pytest.txt
python scripts/hifigan_ncu.py fp32 --time 2>&1 | grep "C=64" > $O/r3b_streamed_64.txt
SS_TF32_PANEL_STREAMED=0 python scripts/hifigan_ncu.py fp32 --time 2>&1 | grep "C=64" > $O/r3b_pertap_64.txt
