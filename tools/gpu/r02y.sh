#!/bin/bash
mkdir -p gpurun_out/r02y
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "rk45" > gpurun_out/r02y/pytest.txt 2>&1; grep -E "RK45|family|passed|failed|Error|assert" gpurun_out/r02y/pytest.txt | head
