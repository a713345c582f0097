#!/bin/bash
# round 2, call V: launch list of the driver's bench command on the final code
mkdir -p gpurun_out/r02V
timeout 600 python bench.py --steps 20 --warmup 5 --no-policy --no-lap --no-large --no-motor-full > gpurun_out/r02V/plain.json 2> gpurun_out/r02V/plain.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02V/launches_bench_steps20.csv python bench.py --steps 20 --warmup 5 --no-policy --no-lap --no-large --no-motor-full > gpurun_out/r02V/ncu.log 2>&1
tail -n 1 gpurun_out/r02V/launches_bench_steps20.csv | cut -c1-200
