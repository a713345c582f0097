#!/bin/bash
mkdir -p gpurun_out/r02E
python tools/ncu_target_dephased.py full > gpurun_out/r02E/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:exo_motor_kernel -s 350 -c 1 -o gpurun_out/r02E/motor_steady python tools/ncu_target_dephased.py full > gpurun_out/r02E/ncu.log 2>&1
tail -1 gpurun_out/r02E/ncu.log
