#!/bin/bash
set -u
O=gpurun_out/r2c9; mkdir -p $O
./scripts/micro/sts_bench > $O/sts_bench.txt 2>&1; cat $O/sts_bench.txt
for v in "0 auto" "4 auto" "0 planes"; do set -- $v
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-cross-check --pack-dense-every $1 --layer-storage $2 > $O/bench_c3_pde$1_$2.json 2> $O/bench_c3_pde$1_$2.err; echo "pde$1 $2 rc=$?" | tee -a $O/rc.txt
python - <<PY
import json
b=json.loads(open("$O/bench_c3_pde$1_$2.json").read().strip().splitlines()[-1]); e=b["e2e"]
print("pde$1 $2: value", round(b["value"],4), "e2e", round(e["value"],4), e["host_transfer"]); print(e["phases_per_step"][-1])
PY
done
