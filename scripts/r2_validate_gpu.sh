#!/bin/bash
# what the driver runs at round end, in one call: GPU tests, smoke, default bench, reference arm (with wall times)
O=gpurun_out
t0=$(date +%s); python -m pytest tests -m gpu -x -q > $O/validate_pytest.txt 2>&1; tail -2 $O/validate_pytest.txt; t1=$(date +%s); echo "pytest wall $((t1-t0)) s"
python -c "import __graft_entry__ as g; g.smoke()" > $O/validate_smoke.txt 2>&1; tail -1 $O/validate_smoke.txt; t2=$(date +%s); echo "smoke wall $((t2-t1)) s"
python bench.py > $O/validate_bench.json 2> $O/validate_bench.err; tail -2 $O/validate_bench.err; t3=$(date +%s); echo "bench wall $((t3-t2)) s"
python bench.py --impl reference --steps 3 --warmup 1 > $O/validate_bench_reference.json 2>&1; t4=$(date +%s); echo "reference arm wall $((t4-t3)) s"
cut -c1-200 $O/validate_bench.json | tail -1
