#!/bin/bash
# Round-2 GPU call 3: fp8 + seed-batching tests, whole GPU suite, C2/C5 bench lines per storage / batching.  Logs -> gpurun_out/r2c3/
set -u
O=gpurun_out/r2c3; mkdir -p $O
export N2V2R_TEST_LOG=$O/testlog
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --debug-modes > $O/probe_tiles_modes.log 2>&1; tail -1 $O/probe_tiles_modes.log
timeout 900 python -m pytest tests/test_gpu_fp8.py -q > $O/pytest_fp8.log 2>&1; echo "fp8 rc=$?" | tee $O/rc.txt
timeout 900 python -m pytest tests/test_gpu_stability.py -q > $O/pytest_stability.log 2>&1; echo "stability rc=$?" | tee -a $O/rc.txt
timeout 1800 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fp8.py --deselect tests/test_gpu_stability.py > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee -a $O/rc.txt
for st in planes e4m3; do
  timeout 600 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e --layer-storage $st > $O/bench_c2_$st.json 2> $O/bench_c2_$st.err; echo "c2 $st rc=$?" | tee -a $O/rc.txt
done
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 1 > $O/bench_c5_seq.json 2> $O/bench_c5_seq.err; echo "c5 seq rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 16 > $O/bench_c5_batch16.json 2> $O/bench_c5_batch16.err; echo "c5 b16 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 1 --layer-storage e4m3 > $O/bench_c5_e4m3.json 2> $O/bench_c5_e4m3.err; echo "c5 e4m3 rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest_fp8.log $O/pytest_stability.log $O/pytest_gpu.log; for f in $O/bench_*.json; do echo $f; cut -c1-400 $f; done
