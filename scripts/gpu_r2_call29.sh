#!/bin/bash
set -u
O=gpurun_out/r2c29; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt; tail -3 $O/pytest_rows.log; grep -E "^E  " $O/pytest_rows.log | head -5
