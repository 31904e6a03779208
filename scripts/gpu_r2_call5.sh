#!/bin/bash
# Round-2 GPU call 5: vectorised converters (tiles), role counters, tile tests.  Logs -> gpurun_out/r2c5/
set -u
O=gpurun_out/r2c5; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tiles.py tests/test_gpu_kernels.py -q > $O/pytest_some.log 2>&1; echo "tests rc=$?" | tee $O/rc.txt
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --debug-modes > $O/probe_tiles_modes.log 2>&1; tail -1 $O/probe_tiles_modes.log
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 > $O/probe_tiles.log 2>&1; tail -1 $O/probe_tiles.log
tail -3 $O/pytest_some.log
