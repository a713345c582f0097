#!/bin/bash
# round 2, call l: full GPU suite (reworked reward checks, shared-stream golden, motor kernel in shared memory) + bench
mkdir -p gpurun_out/r02l
timeout 1800 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02l/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l/pytest_gpu.txt
grep -E "family|shared-stream|options|sweeps/step|passed|failed|rc=|Error|assert" gpurun_out/r02l/pytest_gpu.txt | head -40
timeout 600 python bench.py --steps 20 --warmup 5 --no-policy --no-lap > gpurun_out/r02l/bench_20.json 2> gpurun_out/r02l/bench_20.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r02l/bench_20.json').read().strip().splitlines()[-1])
print('ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'frac %.3f'%d['roofline']['frac'], 'value %.3e'%d['value'])
print(d.get('motor_full')); print(d['roofline_reset'])
PY
