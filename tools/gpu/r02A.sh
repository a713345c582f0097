#!/bin/bash
mkdir -p gpurun_out/r02A
timeout 900 python -m pytest tests/test_gpu_motor.py -m gpu -x -q -s > gpurun_out/r02A/pytest.txt 2>&1; echo "rc=$?" >> gpurun_out/r02A/pytest.txt
grep -E "options|sweeps/step|passed|failed|rc=|Error|assert" gpurun_out/r02A/pytest.txt | head -20
python - <<'PY'
import sys, os, json
sys.path.insert(0, os.getcwd())
import argparse, torch, bench, exo_b200
a = argparse.Namespace(envs_per_gpu=65536, precision="fp32")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
print(json.dumps(bench.bench_motor_full(exo_b200, a, torch.device("cuda:0"), flush)))
PY
