#!/bin/bash
for pdl in 1 0 1 0 1 0; do
  echo "== SS_PDL=$pdl"
  SS_PDL=$pdl timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu-baseline 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | tr '\n' ' '; echo
done
