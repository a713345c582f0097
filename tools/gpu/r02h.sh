#!/bin/bash
# round 2, call h: custom sincos + static actuator indices; parity tests, then the overhead split again
mkdir -p gpurun_out/r02h
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_motor.py -m gpu -x -q > gpurun_out/r02h/pytest.txt 2>&1; tail -3 gpurun_out/r02h/pytest.txt
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02h/$name.json 2> gpurun_out/r02h/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02h/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'frac %.3f'%d['roofline']['frac'], 'value %.3e'%d['value'])
PY
}
run base A=1
run no_gen EXO_DEBUG_NO_GEN=1
run no_gen_no_fused EXO_DEBUG_NO_GEN=1 EXO_DEBUG_NO_FUSED=1
