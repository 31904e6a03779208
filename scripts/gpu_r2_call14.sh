#!/bin/bash
set -u
O=gpurun_out/r2c14; mkdir -p $O
for cfg in "256 4 256" "128 8 256" "192 4 256"; do
  set -- $cfg
  timeout 300 scripts/micro/spmm_proto 50000 0.0264 $1 $2 $3 > $O/spmmB_w$1.log 2>&1; echo "rc=$?" >> $O/spmmB_w$1.log; cat $O/spmmB_w$1.log
done
