#!/bin/bash
mkdir -p gpurun_out/r02z
timeout 1200 python -m pytest tests/test_gpu_motor.py -m gpu -x -q -s -k "reference_integrator" > gpurun_out/r02z/pytest.txt 2>&1; grep -E "rk45|passed|failed|Error|assert" gpurun_out/r02z/pytest.txt | head
