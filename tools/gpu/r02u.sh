#!/bin/bash
# round 2, call u: work-list reservation at load time + next-episode rows staged in shared memory for the fused episode start
mkdir -p gpurun_out/r02u
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_motor.py -m gpu -x -q > gpurun_out/r02u/pytest.txt 2>&1; tail -2 gpurun_out/r02u/pytest.txt
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02u/$name.json 2> gpurun_out/r02u/$name.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02u/$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
except Exception as ex:
    print('$name failed', ex, open('gpurun_out/r02u/$name.err').read()[-600:])
PY
}
run base A=1
run no_gen EXO_DEBUG_NO_GEN=1
run no_gen_no_fused EXO_DEBUG_NO_GEN=1 EXO_DEBUG_NO_FUSED=1
