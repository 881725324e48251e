#!/bin/bash
O=gpurun_out
for pf in 1 0 1 0; do
  echo "== SS_X3_PREFETCH=$pf"
  SS_X3_PREFETCH=$pf timeout 200 python scripts/resunit_one.py 2>&1
  SS_X3_PREFETCH=$pf timeout 300 python scripts/x3_time.py --few 2>&1 | grep "c2"
  SS_X3_PREFETCH=$pf timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline 2>/dev/null | cut -c1-200 | grep -o '"ms_per_step": [0-9.]*'
done
