#!/bin/bash
set -u
O=gpurun_out/r2c23; mkdir -p $O
timeout 900 python scripts/probe_c3_solver.py > $O/probe_c3_solver.log 2>&1; echo "rc=$?" | tee $O/rc.txt; grep -v "^$" $O/probe_c3_solver.log | tail -30 | cut -c1-400
