#!/bin/bash
# round 2, call J: tcgen05 TD7 GEMM with cluster split-k - parity tests, timing per cluster limit beside the default
mkdir -p gpurun_out/r02J
EXO_TD7_GEMM=tc5 timeout 900 python -m pytest tests/test_td7.py -m gpu -q > gpurun_out/r02J/pytest_tc5.txt 2>&1; tail -3 gpurun_out/r02J/pytest_tc5.txt
timeout 600 python tools/td7_probe.py --iters 200 --device-only > gpurun_out/r02J/probe_default.txt 2>&1; grep "device TD7" gpurun_out/r02J/probe_default.txt
for ck in 1 2 4; do
  echo "== tc5 cluster <= $ck"
  EXO_TD7_GEMM=tc5 EXO_TD7_TC5_CLUSTER=$ck timeout 600 python tools/td7_probe.py --iters 200 > gpurun_out/r02J/probe_tc5_ck$ck.txt 2>&1; grep "device TD7\|params after" gpurun_out/r02J/probe_tc5_ck$ck.txt
done
