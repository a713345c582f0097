#!/bin/bash
# round 2, call H: launch lists of the TD7 iteration, mma.sync default and tcgen05 (serialised by ncu; relative durations only)
mkdir -p gpurun_out/r02H
EXO_TD7_GEMM=tc5 timeout 600 python tools/td7_probe.py --iters 2 --device-only > gpurun_out/r02H/probe.txt 2>&1 || exit 1
EXO_TD7_GEMM=tc5 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 600 -c 320 --csv --log-file gpurun_out/r02H/launches_tc5.csv python tools/td7_probe.py --iters 2 --device-only > gpurun_out/r02H/ncu.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 600 -c 320 --csv --log-file gpurun_out/r02H/launches_default.csv python tools/td7_probe.py --iters 2 --device-only > gpurun_out/r02H/ncu2.log 2>&1
tail -2 gpurun_out/r02H/ncu.log
