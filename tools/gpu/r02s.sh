#!/bin/bash
# round 2, call s: episode-start path as one non-inlined copy + a dry pass per SM (instruction-cache warm-up)
mkdir -p gpurun_out/r02s
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02s/pytest.txt 2>&1; tail -2 gpurun_out/r02s/pytest.txt
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02s/$name.json 2> gpurun_out/r02s/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02s/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
PY
}
run warm1 A=1
run warm0 EXO_STEP_WARM=0
run warm2 EXO_STEP_WARM=2
run warm1_no_gen EXO_DEBUG_NO_GEN=1
run warm0_no_gen EXO_DEBUG_NO_GEN=1 EXO_STEP_WARM=0
