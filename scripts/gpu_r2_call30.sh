#!/bin/bash
# Round-2 GPU call 30: wider panels at l = 32 (groups of 4 lanes)
set -u
O=gpurun_out/r2c30; mkdir -p $O
run() { tag=$1; shift; timeout 300 python scripts/probe_rows.py "$@" --no-tiles --rpg 0 > $O/probe_$tag.log 2>&1; echo "$tag rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_$tag.log | grep -o '"list_cap": [0-9]*\|"rows_ms": [0-9.e-]*\|status [0-9]*: .*' | tr '\n' ' '; echo; }
run l32_w256 50000 32 10 20
run l32_w320 50000 32 10 20 --W 320
run l32_w384 50000 32 10 20 --W 384
run l32_w448 50000 32 10 20 --W 448
run l32_w512 50000 32 10 20 --W 512
