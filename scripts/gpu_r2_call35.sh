#!/bin/bash
# Round-2 GPU call 35: pytest -x -m gpu and smoke on the head commit
set -u
O=gpurun_out/r2c35; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "gpu rc=$?" | tee $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -2; grep -E "^FAILED|^E  " $O/pytest_gpu.log | cut -c1-250 | head; tail -1 $O/smoke.log
