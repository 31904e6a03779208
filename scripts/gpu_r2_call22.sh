#!/bin/bash
# Round-2 GPU call 22: K2 projection with two column blocks per A tile read; K chunk variants (time, per-entry error)
set -u
O=gpurun_out/r2c22; mkdir -p $O
timeout 600 python scripts/probe_project.py --variants > $O/probe_project_variants.log 2>&1; echo "variants rc=$?" | tee $O/rc.txt; cat $O/probe_project_variants.log | tail -12
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "project" > $O/pytest_project.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -4 $O/pytest_project.log
