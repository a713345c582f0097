#!/bin/bash
# round 2, call i: staggered load issue by dispatch layer (cycles per layer)
mkdir -p gpurun_out/r02i
B="python bench.py --steps 100 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02i/$name.json 2> gpurun_out/r02i/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02i/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'frac %.3f'%d['roofline']['frac'], 'value %.3e'%d['value'])
PY
}
run s0 EXO_STEP_STAGGER=0
run s300 EXO_STEP_STAGGER=300
run s600 EXO_STEP_STAGGER=600
run s1000 EXO_STEP_STAGGER=1000
run s1500 EXO_STEP_STAGGER=1500
run s2000 EXO_STEP_STAGGER=2000
run s3000 EXO_STEP_STAGGER=3000
