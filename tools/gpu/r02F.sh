#!/bin/bash
# round 2, call F: the TD7 update with its GEMMs on tcgen05 (EXO_TD7_GEMM=tc5) - parity tests, then timing beside the mma.sync default
mkdir -p gpurun_out/r02F
EXO_TD7_GEMM=tc5 timeout 900 python -m pytest tests/test_td7.py -m gpu -q > gpurun_out/r02F/pytest_tc5.txt 2>&1; tail -15 gpurun_out/r02F/pytest_tc5.txt
for m in default tc5; do
  echo "== $m"
  if [ $m = default ]; then E=""; else E="EXO_TD7_GEMM=tc5"; fi
  env $E timeout 600 python tools/td7_probe.py --iters 200 > gpurun_out/r02F/probe_$m.txt 2>&1; tail -6 gpurun_out/r02F/probe_$m.txt
done
