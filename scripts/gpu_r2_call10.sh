#!/bin/bash
# Round-2 GPU call 10: evidence run -- whole suite, smoke, every workload's bench line, l sweep, ncu launch list + full capture.
set -u
O=gpurun_out/r2c10; mkdir -p $O
export N2V2R_TEST_LOG=$O/testlog
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --layer-storage planes --no-cpu > $O/bench_n1_c3_planes.json 2> $O/bench_n1_c3_planes.err; echo "c3 planes rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c3w24 --no-cpu --no-e2e > $O/bench_n1_c3w24_tiles.json 2> $O/bench_n1_c3w24_tiles.err; echo "c3w24 tiles rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c3w24 --no-cpu --no-e2e --layer-storage planes > $O/bench_n1_c3w24_planes.json 2> $O/bench_n1_c3w24_planes.err; echo "c3w24 planes rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 5 > $O/bench_n1_c2.json 2> $O/bench_n1_c2.err; echo "c2 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 5 --no-cpu --no-e2e --layer-storage planes > $O/bench_n1_c2_planes.json 2> $O/bench_n1_c2_planes.err; echo "c2 planes rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c4 > $O/bench_n1_c4.json 2> $O/bench_n1_c4.err; echo "c4 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 > $O/bench_n1_c5.json 2> $O/bench_n1_c5.err; echo "c5 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --seed-batch 1 --no-cpu --no-e2e > $O/bench_n1_c5_seq.json 2> $O/bench_n1_c5_seq.err; echo "c5 seq rc=$?" | tee -a $O/rc.txt
for l in 16 32 64 80; do timeout 200 python scripts/probe_tiles.py 50000 $l 10 10 --only-tiles > $O/probe_tiles_l$l.log 2>&1; tail -1 $O/probe_tiles_l$l.log | cut -c1-300; done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file $O/ncu_launches_c3.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-cross-check > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sparse_mma_kernel -s 4 -c 1 -o $O/ncu_sparse_full python scripts/probe_tiles.py 50000 48 10 2 --only-tiles > $O/ncu_sparse_full.log 2>&1; echo "ncu full rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; echo "ref arm rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; tail -1 $O/smoke.log; for f in $O/bench_*.json; do echo $f; cut -c1-200 $f; done
