#!/bin/bash
# Round-2 GPU call 21: suite + C3 / C2 / C5 lines after the small-dense kernels were specialised on the column-group count.
set -u
O=gpurun_out/r2c21; mkdir -p $O
timeout 2400 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 900 python bench.py --no-cpu --no-e2e > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 5 > $O/bench_n1_c2.json 2> $O/bench_n1_c2.err; echo "c2 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 > $O/bench_n1_c5.json 2> $O/bench_n1_c5.err; echo "c5 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c4 > $O/bench_n1_c4.json 2> $O/bench_n1_c4.err; echo "c4 rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file $O/ncu_launches_c3_rows.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-cross-check > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; grep -E "^FAILED|^E  " $O/pytest_gpu.log | cut -c1-250 | head
python scripts/summarise_launches.py $O/ncu_launches_c3_rows.csv | head -16
for f in $O/bench_*.json; do echo $f; cut -c1-160 $f; done
