#!/bin/bash
# Round-2 GPU call 1: new tile kernel tests, whole GPU suite, probes, short bench lines, ncu, sanitizer.  Logs -> gpurun_out/r2c1/
set -u
O=gpurun_out/r2c1; mkdir -p $O
export N2V2R_TEST_LOG=$O/testlog
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tiles.py -x -q > $O/pytest_tiles.log 2>&1; echo "tiles rc=$?" | tee $O/rc.txt
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 > $O/probe_tiles_c3.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt
timeout 200 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles > $O/probe_tiles_c3_only.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_tiles.py > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e > $O/bench_c3_tiles.json 2> $O/bench_c3_tiles.err; echo "bench tiles rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --layer-storage planes > $O/bench_c3_planes.json 2> $O/bench_c3_planes.err; echo "bench planes rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sparse_mma_kernel -s 4 -c 1 -o $O/ncu_sparse_full python scripts/probe_tiles.py 50000 48 10 2 --only-tiles > $O/ncu_sparse_full.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt
timeout 600 compute-sanitizer --tool memcheck python scripts/probe_tiles.py 3000 48 10 1 --only-tiles > $O/sanitizer_memcheck_tiles.log 2>&1; echo "memcheck rc=$?" | tee -a $O/rc.txt
timeout 600 compute-sanitizer --tool racecheck python scripts/probe_tiles.py 3000 48 10 1 --only-tiles > $O/sanitizer_racecheck_tiles.log 2>&1; echo "racecheck rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_tiles.log $O/pytest_gpu.log; cat $O/probe_tiles_c3.log | tail -2; cat $O/bench_c3_tiles.json | cut -c1-1500
