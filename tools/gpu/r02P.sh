#!/bin/bash
# round 2, call P: ncu --set full of steady-state tcgen05 TD7 GEMM launches + the launch list of ~2 iterations
mkdir -p gpurun_out/r02P
timeout 300 python tools/td7_probe.py --iters 2 --device-only > gpurun_out/r02P/plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc5_kernel -s 600 -c 12 -o gpurun_out/r02P/td7_tc5 python tools/td7_probe.py --iters 2 --device-only > gpurun_out/r02P/ncu.log 2>&1
tail -n 2 gpurun_out/r02P/ncu.log
