#!/bin/bash
# round 2, call C: exo_motor_kernel compiled for 4/5/6/7 resident blocks per SM
mkdir -p gpurun_out/r02C
timeout 600 python -m pytest tests/test_gpu_motor.py -m gpu -x -q > gpurun_out/r02C/pytest.txt 2>&1; tail -1 gpurun_out/r02C/pytest.txt
for v in default mb7; do
  if [ $v = default ]; then L=""; else L="EXO_LIB=$PWD/_variants/libexo_$v.so"; fi
  env $L python - <<'PY'
import sys, os, json, argparse, torch
sys.path.insert(0, os.getcwd())
import bench, exo_b200
a = argparse.Namespace(envs_per_gpu=65536, precision="fp32")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
r = bench.bench_motor_full(exo_b200, a, torch.device("cuda:0"), flush)
print(os.environ.get("EXO_LIB", "default"), "ms_per_step %.4f value %.3e" % (r["ms_per_step"], r["value"]))
PY
done
