#!/bin/bash
# round 2, call R: full GPU suite on the cleaned-up episode-start code
mkdir -p gpurun_out/r02R
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02R/pytest_gpu.txt 2>&1; tail -n 3 gpurun_out/r02R/pytest_gpu.txt
