#!/bin/bash
# Round-2 two-GPU call: invariance tests (layer sharding, row slabs), C3 on 2 GPUs.  Logs -> gpurun_out/r2m2/
set -u
O=gpurun_out/r2m2; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > $O/pytest_multi.log 2>&1; echo "multi rc=$?" | tee $O/rc.txt; tail -15 $O/pytest_multi.log
cp gpurun_out/multi_gpu_invariance.json $O/ 2>/dev/null
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 scripts/multi_gpu_invariance.py c3 20000 $O/invariance_c3_n20000_w2.json > $O/inv_c3.log 2>&1; echo "inv c3 rc=$?" | tee -a $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 scripts/multi_gpu_invariance.py c2 20000 $O/invariance_c2_onelayer_w2.json --one-layer > $O/inv_slab.log 2>&1; echo "inv slab rc=$?" | tee -a $O/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_c3_n2.json 2> $O/bench_c3_n2.err; echo "bench n2 rc=$?" | tee -a $O/rc.txt
tail -2 $O/inv_c3.log | cut -c1-900; tail -2 $O/inv_slab.log | cut -c1-900; cut -c1-400 $O/bench_c3_n2.json; tail -5 $O/bench_c3_n2.err
