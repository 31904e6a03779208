#!/bin/bash
# Round-2 GPU call 17: the no-allocation entry point's test alone in a fresh process, then the whole suite with -x (as the round's end runs it).
set -u
O=gpurun_out/r2c17; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" | tee $O/rc.txt; tail -4 $O/pytest_no_alloc_alone.log
export N2V2R_TEST_LOG=$O/testlog
timeout 2400 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; grep -E "^FAILED|^E  " $O/pytest_gpu.log | cut -c1-250 | head -20
