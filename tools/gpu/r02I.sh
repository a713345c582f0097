#!/bin/bash
# round 2, call I: warp-specialised tcgen05 TD7 GEMM - timing beside the default
mkdir -p gpurun_out/r02I
for m in default tc5; do
  if [ $m = default ]; then E=""; else E="EXO_TD7_GEMM=tc5"; fi
  env $E timeout 600 python tools/td7_probe.py --iters 200 --device-only > gpurun_out/r02I/probe_$m.txt 2>&1; grep "device TD7" gpurun_out/r02I/probe_$m.txt
done
