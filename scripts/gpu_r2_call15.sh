#!/bin/bash
set -u
O=gpurun_out/r2c15; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt; tail -5 $O/pytest_rows.log
timeout 300 python scripts/probe_rows.py 50000 48 10 20 --no-tiles > $O/probe_rows_c3.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt; tail -2 $O/probe_rows_c3.log | cut -c1-1200
timeout 300 python scripts/probe_rows.py 50000 48 10 20 --no-tiles --W 128 > $O/probe_rows_c3_w128.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt; tail -2 $O/probe_rows_c3_w128.log | cut -c1-1200
timeout 300 python scripts/probe_rows.py 50000 48 10 20 --no-tiles --W 320 > $O/probe_rows_c3_w320.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt; tail -2 $O/probe_rows_c3_w320.log | cut -c1-1200
timeout 300 python scripts/probe_rows.py 50000 80 10 20 --no-tiles > $O/probe_rows_l80.log 2>&1; echo "probe l80 rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_rows_l80.log | cut -c1-900
timeout 600 python bench.py --steps 3 --warmup 3 > $O/bench_c3_rows.json 2> $O/bench_c3_rows.err; echo "bench rc=$?" | tee -a $O/rc.txt; cut -c1-400 $O/bench_c3_rows.json; tail -3 $O/bench_c3_rows.err
