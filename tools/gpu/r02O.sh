#!/bin/bash
# round 2, call O: DRAM traffic of exo_step_kernel under the bench's own flush (dirty L2) and under a clean cold L2
mkdir -p gpurun_out/r02O
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum
timeout 300 python tools/ncu_target_dirty_l2.py write > gpurun_out/r02O/plain.log 2>&1 || exit 1
timeout 600 ncu --cache-control none --clock-control none -k regex:exo_step_kernel -s 5 -c 12 --metrics $M --csv --log-file gpurun_out/r02O/dirty_l2.csv python tools/ncu_target_dirty_l2.py write > gpurun_out/r02O/ncu1.log 2>&1
timeout 600 ncu --cache-control none --clock-control none -k regex:exo_step_kernel -s 5 -c 12 --metrics $M --csv --log-file gpurun_out/r02O/clean_l2.csv python tools/ncu_target_dirty_l2.py read > gpurun_out/r02O/ncu2.log 2>&1
tail -2 gpurun_out/r02O/ncu1.log gpurun_out/r02O/ncu2.log
