#!/bin/bash
# round 2, call b: GPU tests (new: EXO_MOTOR_FULL, two-shard independence) + the de-phased bench
mkdir -p gpurun_out/r02b
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02b/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b/pytest_gpu.txt
tail -25 gpurun_out/r02b/pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02b/bench_20.json 2> gpurun_out/r02b/bench_20.err; echo "bench rc=$?"
tail -3 gpurun_out/r02b/bench_20.err
