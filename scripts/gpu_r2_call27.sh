#!/bin/bash
# Round-2 GPU call 27: validation of the head commit as the driver runs it: pytest -x -m gpu, smoke, default bench, reference arm.
set -u
O=gpurun_out/r2c27; mkdir -p $O
timeout 2400 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; echo "ref arm rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 5 --no-cpu > $O/bench_n1_c2.json 2> $O/bench_n1_c2.err; echo "c2 rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; grep -E "^FAILED|^E  " $O/pytest_gpu.log | cut -c1-250 | head; tail -1 $O/smoke.log
for f in $O/bench_*.json; do echo $f; cut -c1-200 $f; done
