#!/bin/bash
# TS kernel (A tile in tensor memory): launch time against the operand block width l
set -u
O=gpurun_out/r2c12; mkdir -p $O
for l in 16 32 64 80; do
  timeout 300 python scripts/probe_tiles.py 50000 $l 10 10 --only-tiles --ts > $O/probe_ts_l$l.log 2>&1; echo "l=$l rc=$?" | tee -a $O/rc.txt; tail -1 $O/probe_ts_l$l.log | cut -c1-700
done
