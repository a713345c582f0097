#!/bin/bash
# round 2, call Z: the driver's sequence once more on the committed head: GPU tests (-x), smoke, default bench, reference arm
mkdir -p gpurun_out/r02Z
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02Z/pytest_gpu.txt 2>&1; tail -n 2 gpurun_out/r02Z/pytest_gpu.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02Z/smoke.txt 2>&1; tail -n 1 gpurun_out/r02Z/smoke.txt
timeout 900 python bench.py > gpurun_out/r02Z/bench_default.json 2> gpurun_out/r02Z/bench_default.err; echo "default rc $?"
timeout 900 python bench.py --impl reference > gpurun_out/r02Z/bench_reference.json 2> gpurun_out/r02Z/bench_reference.err; echo "ref rc $?"
