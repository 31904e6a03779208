#!/bin/bash
set -u
O=gpurun_out/r2c11; mkdir -p $O
timeout 120 python scripts/probe_tiles.py 3000 48 10 3 --only-tiles --ts > $O/probe_ts_small.log 2>&1; echo "small rc=$?" | tee $O/rc.txt; tail -2 $O/probe_ts_small.log | cut -c1-900
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --ts > $O/probe_ts_c3.log 2>&1; echo "c3 rc=$?" | tee -a $O/rc.txt; tail -2 $O/probe_ts_c3.log | cut -c1-900
