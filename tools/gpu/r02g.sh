#!/bin/bash
# round 2, call g: split the steady-state overhead: fused begin path vs generation passes
mkdir -p gpurun_out/r02g
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02g/$name.json 2> gpurun_out/r02g/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02g/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
PY
}
run base A=1
run no_gen EXO_DEBUG_NO_GEN=1
run no_fused EXO_DEBUG_NO_FUSED=1
run no_gen_no_fused EXO_DEBUG_NO_GEN=1 EXO_DEBUG_NO_FUSED=1
B="$B --no-flush"
run noflush A=1
run noflush_no_gen EXO_DEBUG_NO_GEN=1
