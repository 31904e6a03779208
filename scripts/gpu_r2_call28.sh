#!/bin/bash
# Round-2 GPU call 28: groups of 8 lanes for l = 32 / 64
set -u
O=gpurun_out/r2c28; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee $O/rc.txt; tail -3 $O/pytest_rows.log
run() { tag=$1; shift; timeout 300 python scripts/probe_rows.py "$@" --no-tiles --rpg 0 > $O/probe_$tag.log 2>&1; echo "$tag rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_$tag.log | grep -o '"rows_ms": [0-9.e-]*'; }
run l32_lanes8 50000 32 10 20
run l32_lanes4 50000 32 10 20 --lanes4
run l64_lanes8 50000 64 10 20
run l64_lanes4 50000 64 10 20 --lanes4
run l32_lanes8_w320 50000 32 10 20 --W 320
run l32_lanes8_w384 50000 32 10 20 --W 384
