#!/bin/bash
# Round-2 GPU call 4: role cycle counters of the sparse kernel, e4m3 after the colmax fix, seed-batch sizes.  Logs -> gpurun_out/r2c4/
set -u
O=gpurun_out/r2c4; mkdir -p $O
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --debug-modes > $O/probe_tiles_modes.log 2>&1; tail -1 $O/probe_tiles_modes.log
timeout 600 python -m pytest tests/test_gpu_fp8.py tests/test_gpu_tiles.py tests/test_gpu_kernels.py -q > $O/pytest_some.log 2>&1; echo "tests rc=$?" | tee $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu --no-e2e --layer-storage e4m3 > $O/bench_c2_e4m3.json 2> $O/bench_c2_e4m3.err; echo "c2 e4m3 rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch 1 --layer-storage e4m3 > $O/bench_c5_e4m3.json 2> $O/bench_c5_e4m3.err; echo "c5 e4m3 rc=$?" | tee -a $O/rc.txt
for sb in 2 3 6; do
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu --no-e2e --seed-batch $sb > $O/bench_c5_batch$sb.json 2> $O/bench_c5_batch$sb.err; echo "c5 b$sb rc=$?" | tee -a $O/rc.txt
done
tail -3 $O/pytest_some.log; for f in $O/bench_*.json; do echo $f; cut -c1-200 $f; done
