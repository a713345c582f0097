#!/bin/bash
mkdir -p gpurun_out/r02v
python tools/ncu_target_dephased.py > gpurun_out/r02v/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:exo_step_kernel -s 350 -c 1 -o gpurun_out/r02v/step_dephased python tools/ncu_target_dephased.py > gpurun_out/r02v/ncu.log 2>&1
tail -2 gpurun_out/r02v/ncu.log; ls -la gpurun_out/r02v
