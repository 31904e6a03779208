#!/bin/bash
set -u
O=gpurun_out/r2c19; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt; tail -5 $O/pytest_rows.log
for l in 48 80 16; do
  timeout 300 python scripts/probe_rows.py 50000 $l 10 20 --no-tiles --rpg 0 > $O/probe_rows_l$l.log 2>&1; echo "l=$l rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_rows_l$l.log | cut -c150-600
done
