#!/bin/bash
# Round-2 eight-GPU call: scaling lines of C3 (N = 8, 4), C2 with row slabs, C5 with seed sharding.  Logs -> gpurun_out/r2m8/
set -u
O=gpurun_out/r2m8; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1
run() { # n port args...
  local n=$1; local port=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@"
}
run 8 29541 --steps 5 --warmup 3 --no-cpu > $O/bench_c3_n8.json 2> $O/bench_c3_n8.err; echo "c3 n8 rc=$?" | tee $O/rc.txt
run 4 29542 --steps 5 --warmup 3 --no-cpu > $O/bench_c3_n4.json 2> $O/bench_c3_n4.err; echo "c3 n4 rc=$?" | tee -a $O/rc.txt
run 8 29543 --steps 5 --warmup 3 --no-cpu --workload c2 > $O/bench_c2_n8_slabs.json 2> $O/bench_c2_n8_slabs.err; echo "c2 n8 rc=$?" | tee -a $O/rc.txt
run 8 29544 --steps 3 --warmup 2 --no-cpu --workload c5 > $O/bench_c5_n8.json 2> $O/bench_c5_n8.err; echo "c5 n8 rc=$?" | tee -a $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29545 scripts/multi_gpu_invariance.py c2 20000 $O/invariance_c2_w8_slabs.json > $O/inv_c2_w8.log 2>&1; echo "inv c2 w8 rc=$?" | tee -a $O/rc.txt
for f in $O/bench_*.json; do echo $f; cut -c1-300 $f; done; tail -2 $O/inv_c2_w8.log | cut -c1-600
