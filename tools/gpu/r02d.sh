#!/bin/bash
# round 2, call d: where do the 10 us between the step kernel (21 us) and the steady-state step (32 us) go?
mkdir -p gpurun_out/r02d
B="python bench.py --steps 200 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02d/$name.json 2> gpurun_out/r02d/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02d/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
PY
}
run default A=1
run warps4 EXO_GEN_WARPS=4
run warps2 EXO_GEN_WARPS=2
run noprio EXO_GEN_NOPRIO=1
run grid148 EXO_GEN_GRID=148
run grid296 EXO_GEN_GRID=296
B="$B --no-dephase"; run nodephase A=1
