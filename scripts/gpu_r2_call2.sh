#!/bin/bash
# Round-2 GPU call 2: tile tests (all), loader tests, tile probe, light ncu of the sparse kernel.  Logs -> gpurun_out/r2c2/
set -u
O=gpurun_out/r2c2; mkdir -p $O
export N2V2R_TEST_LOG=$O/testlog
timeout 900 python -m pytest tests/test_gpu_tiles.py tests/test_gpu_loader.py -q > $O/pytest_tiles_loader.log 2>&1; echo "tiles+loader rc=$?" | tee $O/rc.txt
timeout 300 python scripts/probe_tiles.py 50000 48 10 20 > $O/probe_tiles_c3.log 2>&1; echo "probe rc=$?" | tee -a $O/rc.txt
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,sm__cycles_elapsed.avg,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum
timeout 300 ncu --metrics $M --clock-control none -k regex:sparse_mma_kernel -s 4 -c 2 --csv --log-file $O/ncu_sparse_light.csv python scripts/probe_tiles.py 50000 48 10 2 --only-tiles > $O/ncu_sparse_light.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt
tail -5 $O/pytest_tiles_loader.log; tail -1 $O/probe_tiles_c3.log; grep -v "^==" $O/ncu_sparse_light.csv | cut -d, -f5,13- | tail -24
