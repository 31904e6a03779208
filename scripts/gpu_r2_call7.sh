#!/bin/bash
# Round-2 GPU call 7: the CTA-pair sparse kernel.  Logs -> gpurun_out/r2c7/
set -u
O=gpurun_out/r2c7; mkdir -p $O
timeout 120 python scripts/probe_tiles.py 3000 48 10 3 --only-tiles --cta2 > $O/probe_cta2_small.log 2>&1; echo "small rc=$?" | tee $O/rc.txt; tail -2 $O/probe_cta2_small.log | cut -c1-600
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --cta2 > $O/probe_cta2_c3.log 2>&1; echo "c3 rc=$?" | tee -a $O/rc.txt; tail -2 $O/probe_cta2_c3.log | cut -c1-600
