#!/bin/bash
# round 2, call r: the driver's sequence on the current code: GPU tests, smoke, default bench, reference arm, launch list
mkdir -p gpurun_out/r02r
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02r/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02r/pytest_gpu.txt; tail -3 gpurun_out/r02r/pytest_gpu.txt
timeout 600 python __graft_entry__.py smoke > gpurun_out/r02r/smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02r/smoke.txt
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02r/bench_steps20.json 2> gpurun_out/r02r/bench_steps20.err; echo "bench rc=$?"
timeout 900 python bench.py > gpurun_out/r02r/bench_default.json 2> gpurun_out/r02r/bench_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02r/bench_reference.json 2> gpurun_out/r02r/bench_reference.err; echo "ref rc=$?"
python bench.py --steps 20 --warmup 5 --no-policy --no-lap --no-large --no-motor-full > gpurun_out/r02r/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02r/launches_bench_steps20.csv python bench.py --steps 20 --warmup 5 --no-policy --no-lap --no-large --no-motor-full > gpurun_out/r02r/ncu.log 2>&1
python - <<'PY'
import json
for f in ('bench_steps20','bench_default','bench_reference'):
    d=json.loads(open('gpurun_out/r02r/%s.json'%f).read().strip().splitlines()[-1])
    print(f, 'value %.3e'%d['value'], 'ms %.4f'%d['ms_per_step'], d.get('roofline',{}).get('frac'), 'e2e %.3e'%d['e2e']['value'])
PY
