#!/bin/bash
# Round-2 GPU call 6: whole suite, C3 default line with e2e (hybrid transfer), C2/C5 variants after the collector change.  Logs -> gpurun_out/r2c6/
set -u
O=gpurun_out/r2c6; mkdir -p $O
export N2V2R_TEST_LOG=$O/testlog
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 > $O/probe_tiles.log 2>&1; tail -1 $O/probe_tiles.log
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > $O/bench_c3.json 2> $O/bench_c3.err; echo "c3 rc=$?" | tee -a $O/rc.txt
for st in planes e4m3; do
  timeout 600 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e --layer-storage $st > $O/bench_c2_$st.json 2> $O/bench_c2_$st.err; echo "c2 $st rc=$?" | tee -a $O/rc.txt
done
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 1 > $O/bench_c5_seq.json 2> $O/bench_c5_seq.err; echo "c5 seq rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 3 > $O/bench_c5_batch3.json 2> $O/bench_c5_batch3.err; echo "c5 b3 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 6 > $O/bench_c5_batch6.json 2> $O/bench_c5_batch6.err; echo "c5 b6 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 1 --layer-storage e4m3 > $O/bench_c5_e4m3.json 2> $O/bench_c5_e4m3.err; echo "c5 e4m3 rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -3; for f in $O/bench_*.json; do echo $f; cut -c1-160 $f; done
