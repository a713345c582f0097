#!/bin/bash
mkdir -p gpurun_out/r02q
B="python bench.py --steps 100 --warmup 20 --no-policy --no-lap --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02q/$name.json 2> gpurun_out/r02q/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02q/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'frac %.3f'%d['roofline']['frac'], 'value %.3e'%d['value'], 'large', d['roofline'].get('large_population',{}).get('kernel_ms'), d['roofline'].get('large_population',{}).get('frac'))
PY
}
run s1470 A=1
run s0 EXO_STEP_STAGGER=0
run s1000 EXO_STEP_STAGGER=1000
run s2000 EXO_STEP_STAGGER=2000
python tools/step_timeline.py > gpurun_out/r02q/timeline.txt 2>&1; cat gpurun_out/r02q/timeline.txt
