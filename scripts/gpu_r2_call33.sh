#!/bin/bash
# Round-2 GPU call 33: validation of the head commit: pytest -x -m gpu, smoke, default bench (pinned host graphs after the timed steps), 24-bit weights
set -u
O=gpurun_out/r2c33; mkdir -p $O
timeout 2400 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --steps 5 > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c3w24 --no-cpu --no-e2e --steps 5 > $O/bench_n1_c3w24_rows.json 2> $O/bench_n1_c3w24_rows.err; echo "c3w24 rows rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; grep -E "^FAILED|^E  " $O/pytest_gpu.log | cut -c1-250 | head; tail -1 $O/smoke.log
for f in $O/bench_*.json; do echo $f; cut -c1-160 $f; done; tail -3 $O/bench_n1_c3.err
