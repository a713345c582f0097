#!/bin/bash
# round 2, call K: tcgen05 GEMM as the TD7 default - full GPU suite, td7 tests in the other GEMM modes, probe, bench
mkdir -p gpurun_out/r02K
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02K/pytest_gpu.txt 2>&1; tail -3 gpurun_out/r02K/pytest_gpu.txt
EXO_TD7_GEMM=mma timeout 600 python -m pytest tests/test_td7.py -m gpu -q > gpurun_out/r02K/pytest_td7_mma.txt 2>&1; tail -1 gpurun_out/r02K/pytest_td7_mma.txt
EXO_TD7_TC5_CLUSTER=4 timeout 600 python -m pytest tests/test_td7.py -m gpu -q > gpurun_out/r02K/pytest_td7_cluster4.txt 2>&1; tail -1 gpurun_out/r02K/pytest_td7_cluster4.txt
timeout 600 python tools/td7_probe.py --iters 200 > gpurun_out/r02K/probe.txt 2>&1; grep "device TD7\|params after\|torch" gpurun_out/r02K/probe.txt
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02K/bench.json 2> gpurun_out/r02K/bench.err; tail -c 600 gpurun_out/r02K/bench.json
