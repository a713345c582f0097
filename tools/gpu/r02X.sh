#!/bin/bash
# round 2, call X: BASELINE config 5 end to end on the final code (TD7 update on tcgen05)
mkdir -p gpurun_out/r02X
timeout 600 python tools/train_probe.py --replicas 1 64 1024 --episodes 8 > gpurun_out/r02X/train_probe.txt 2>&1; tail -n 4 gpurun_out/r02X/train_probe.txt
EXO_TD7_GEMM=mma timeout 600 python tools/train_probe.py --replicas 1024 --episodes 8 > gpurun_out/r02X/train_probe_mma.txt 2>&1; tail -n 1 gpurun_out/r02X/train_probe_mma.txt
