#!/bin/bash
set -u
O=gpurun_out/r2c8; mkdir -p $O
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 --only-tiles --cta2 > $O/probe_cta2_c3.log 2>&1; echo "c3 rc=$?" | tee $O/rc.txt; tail -1 $O/probe_cta2_c3.log | cut -c1-3000
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu --no-cross-check > $O/bench_c3.json 2> $O/bench_c3.err; echo "c3 bench rc=$?" | tee -a $O/rc.txt; cut -c1-300 $O/bench_c3.json
