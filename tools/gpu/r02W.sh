#!/bin/bash
# round 2, call W: ncu --set full of one steady-state exo_step_kernel launch (final code, de-phased population)
mkdir -p gpurun_out/r02W
timeout 600 python tools/ncu_target_dephased.py > gpurun_out/r02W/plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:exo_step_kernel -s 350 -c 1 -o gpurun_out/r02W/step_steady python tools/ncu_target_dephased.py > gpurun_out/r02W/ncu.log 2>&1
tail -n 2 gpurun_out/r02W/ncu.log
