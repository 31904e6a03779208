#!/bin/bash
# C3 on N GPUs with the row-list storage (N = $1).  Logs -> gpurun_out/r2mr$1/
set -u
N=$1
O=gpurun_out/r2mr$N; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N bench.py --gpus $N --steps 5 --warmup 3 --no-cpu > $O/bench_c3_n$N.json 2> $O/bench_c3_n$N.err; echo "c3 n$N rc=$?" | tee $O/rc.txt
cut -c1-300 $O/bench_c3_n$N.json; tail -3 $O/bench_c3_n$N.err
