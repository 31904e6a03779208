#!/bin/bash
# Round-2 GPU call 18: ncu evidence for the row-list path (launch list of a bench run, one full capture of rows_product_kernel),
# the reference arm and the remaining one-GPU bench lines on the current library.
set -u
O=gpurun_out/r2c18; mkdir -p $O
timeout 900 python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-cross-check > $O/bench_plain.json 2> $O/bench_plain.err; echo "plain rc=$?" | tee $O/rc.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file $O/ncu_launches_c3_rows.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-cross-check > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?" | tee -a $O/rc.txt
timeout 300 python scripts/probe_rows.py 50000 48 10 2 --no-tiles --rpg 0 > $O/probe_rows_plain.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rows_product_kernel -s 4 -c 1 -o $O/ncu_rows_full python scripts/probe_rows.py 50000 48 10 2 --no-tiles --rpg 0 > $O/ncu_rows_full.log 2>&1; echo "ncu full rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --layer-storage tiles --no-cpu --no-e2e > $O/bench_n1_c3_tiles.json 2> $O/bench_n1_c3_tiles.err; echo "c3 tiles rc=$?" | tee -a $O/rc.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; echo "ref arm rc=$?" | tee -a $O/rc.txt
python scripts/summarise_launches.py $O/ncu_launches_c3_rows.csv | head -20
cat $O/rc.txt
