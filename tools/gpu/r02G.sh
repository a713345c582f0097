#!/bin/bash
# round 2, call G: tcgen05 TD7 GEMM with register prefetch - parity subset + timing
mkdir -p gpurun_out/r02G
EXO_TD7_GEMM=tc5 timeout 900 python -m pytest tests/test_td7.py -m gpu -q > gpurun_out/r02G/pytest_tc5.txt 2>&1; tail -3 gpurun_out/r02G/pytest_tc5.txt
for m in default tc5; do
  echo "== $m"
  if [ $m = default ]; then E=""; else E="EXO_TD7_GEMM=tc5"; fi
  env $E timeout 600 python tools/td7_probe.py --iters 200 > gpurun_out/r02G/probe_$m.txt 2>&1; grep "device TD7" gpurun_out/r02G/probe_$m.txt
  env $E EXO_TD7_GRAPH=0 timeout 600 python tools/td7_probe.py --iters 200 > gpurun_out/r02G/probe_nograph_$m.txt 2>&1; grep "device TD7" gpurun_out/r02G/probe_nograph_$m.txt
done
