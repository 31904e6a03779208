#!/bin/bash
# Round-2 GPU call 31: row-list layers built for the solver's block width (512-column panels at l = 32): rows tests, C3 bench
set -u
O=gpurun_out/r2c31; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt; tail -3 $O/pytest_rows.log; grep -E "^E  " $O/pytest_rows.log | head -5
timeout 900 python bench.py --steps 5 > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee -a $O/rc.txt; cut -c1-200 $O/bench_n1_c3.json; tail -3 $O/bench_n1_c3.err
