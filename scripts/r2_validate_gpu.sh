#!/bin/bash
# what the driver runs at round end, in one call: GPU tests, smoke, default bench, reference arm
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/validate_pytest.txt 2>&1; tail -4 $O/validate_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/validate_smoke.txt 2>&1; tail -2 $O/validate_smoke.txt
python bench.py > $O/validate_bench.json 2> $O/validate_bench.err; tail -2 $O/validate_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/validate_bench_reference.json 2>&1
