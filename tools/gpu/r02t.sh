#!/bin/bash
# round 2, call t: what makes the fused episode start cost ~6 us after an L2 flush? variants without its loads / its stores
mkdir -p gpurun_out/r02t
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02t/$name.json 2> gpurun_out/r02t/$name.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02t/$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
except Exception as ex:
    print('$name failed', ex, open('gpurun_out/r02t/$name.err').read()[-600:])
PY
}
run base_no_gen EXO_DEBUG_NO_GEN=1
run noload_no_gen EXO_DEBUG_NO_GEN=1 EXO_LIB=$PWD/_variants/libexo_noload.so
run nostore_no_gen EXO_DEBUG_NO_GEN=1 EXO_LIB=$PWD/_variants/libexo_nostore.so
run base_no_gen_nostagger EXO_DEBUG_NO_GEN=1 EXO_STEP_STAGGER=0
