#!/bin/bash
# Round-2 GPU call 34: the default bench line on the head commit
set -u
O=gpurun_out/r2c34; mkdir -p $O
timeout 900 python bench.py --steps 5 > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee $O/rc.txt
cut -c1-160 $O/bench_n1_c3.json; tail -3 $O/bench_n1_c3.err
