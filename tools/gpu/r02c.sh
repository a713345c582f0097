#!/bin/bash
# round 2, call c: episode prefetch (double-buffered parameters, fused begin) - tests + bench
mkdir -p gpurun_out/r02c
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02c/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02c/pytest_gpu.txt
tail -15 gpurun_out/r02c/pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-policy --no-lap > gpurun_out/r02c/bench_20.json 2> gpurun_out/r02c/bench_20.err; echo "bench rc=$?"
tail -3 gpurun_out/r02c/bench_20.err
timeout 600 python bench.py --steps 1000 --warmup 50 --no-policy --no-lap --no-large --no-motor-full > gpurun_out/r02c/bench_1000.json 2> gpurun_out/r02c/bench_1000.err; echo "bench rc=$?"
timeout 300 python tools/reset_probe.py > gpurun_out/r02c/reset_probe.txt 2>&1
