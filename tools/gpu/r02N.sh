#!/bin/bash
# round 2, call N: the allocation-free policy entry point - its test alone in a fresh process, then the whole GPU suite (-x, as at round end)
mkdir -p gpurun_out/r02N
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02N/pytest_noalloc_alone.txt 2>&1; tail -3 gpurun_out/r02N/pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02N/pytest_gpu.txt 2>&1; tail -3 gpurun_out/r02N/pytest_gpu.txt
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02N/sanitizer.txt 2>&1; tail -4 gpurun_out/r02N/sanitizer.txt
