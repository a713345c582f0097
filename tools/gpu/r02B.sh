#!/bin/bash
mkdir -p gpurun_out/r02B
python tools/ncu_target.py 65536 3 full > gpurun_out/r02B/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:exo_motor_kernel -s 1 -c 1 -o gpurun_out/r02B/motor python tools/ncu_target.py 65536 3 full > gpurun_out/r02B/ncu.log 2>&1
tail -2 gpurun_out/r02B/ncu.log
