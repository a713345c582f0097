#!/bin/bash
# round 2, call Q: driver-like sequence on the current code: smoke, default bench, --steps 20 --warmup 5 bench, reference arm
mkdir -p gpurun_out/r02Q
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02Q/smoke.txt 2>&1; tail -n 2 gpurun_out/r02Q/smoke.txt
timeout 900 python bench.py > gpurun_out/r02Q/bench_default.json 2> gpurun_out/r02Q/bench_default.err; echo "default rc $?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02Q/bench_20_5.json 2> gpurun_out/r02Q/bench_20_5.err; echo "20/5 rc $?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02Q/bench_reference.json 2> gpurun_out/r02Q/bench_reference.err; echo "ref rc $?"; tail -c 400 gpurun_out/r02Q/bench_reference.json
