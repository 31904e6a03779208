#!/bin/bash
# Round-2 GPU call 32: evidence with 512-column panels: default bench, 24-bit weights, ncu launch list + full capture
set -u
O=gpurun_out/r2c32; mkdir -p $O
timeout 900 python bench.py --steps 5 > $O/bench_n1_c3.json 2> $O/bench_n1_c3.err; echo "c3 rc=$?" | tee $O/rc.txt
timeout 900 python bench.py --workload c3w24 --no-cpu --no-e2e --steps 5 > $O/bench_n1_c3w24_rows.json 2> $O/bench_n1_c3w24_rows.err; echo "c3w24 rows rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file $O/ncu_launches_c3_rows.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-cross-check > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rows_product_kernel -s 4 -c 1 -o $O/ncu_rows_full_l32 python scripts/probe_rows.py 50000 32 10 2 --no-tiles --rpg 0 --W 512 > $O/ncu_rows_full.log 2>&1; echo "ncu full rc=$?" | tee -a $O/rc.txt
python scripts/summarise_launches.py $O/ncu_launches_c3_rows.csv | head -8
for f in $O/bench_*.json; do echo $f; cut -c1-160 $f; done
