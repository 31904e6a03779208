#!/bin/bash
# Round-2 GPU call 26: row blocks of half the size, two CTAs per SM (one row per lane group)
set -u
O=gpurun_out/r2c26; mkdir -p $O
run() { tag=$1; shift; timeout 300 python scripts/probe_rows.py "$@" --no-tiles --rpg 0 > $O/probe_$tag.log 2>&1; echo "$tag rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_$tag.log | cut -c100-520; }
run l32_default 50000 32 10 20
run l32_rb169 50000 32 10 20 --rb 169
run l32_rb169_w192 50000 32 10 20 --rb 169 --W 192
run l32_rb169_1persm 50000 32 10 20 --rb 169 --one-per-sm
run l48_rb169_w128 50000 48 10 20 --rb 169 --W 128
run l16_rb169 50000 16 10 20 --rb 169
timeout 600 python -m pytest tests/test_gpu_rows.py -x -q -m gpu > $O/pytest_rows.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_rows.log
