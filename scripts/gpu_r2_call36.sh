#!/bin/bash
# Round-2 GPU call 36: ncu --set full of the projection's SYRK launch (stream_mma_kernel, two column blocks per A tile)
set -u
O=gpurun_out/r2c36; mkdir -p $O
timeout 400 ncu --set full --clock-control none --import-source on -k regex:stream_mma_kernel -s 1 -c 1 -o $O/ncu_project_full python scripts/probe_project.py > $O/ncu_project.log 2>&1; echo "ncu rc=$?" | tee $O/rc.txt; tail -3 $O/ncu_project.log | cut -c1-200
