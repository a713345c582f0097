#!/bin/bash
# round 2, call x: warps holding an env whose episode ends skip the stagger wait (head start hides the episode-start tail)
mkdir -p gpurun_out/r02x
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02x/pytest.txt 2>&1; tail -2 gpurun_out/r02x/pytest.txt
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02x/$name.json 2> gpurun_out/r02x/$name.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02x/$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'frac %.3f'%d['roofline']['frac'], 'value %.3e'%d['value'])
except Exception as ex:
    print('$name failed', ex, open('gpurun_out/r02x/$name.err').read()[-600:])
PY
}
run s1470 A=1
run s1000 EXO_STEP_STAGGER=1000
run s2000 EXO_STEP_STAGGER=2000
run s2500 EXO_STEP_STAGGER=2500
run s1470_no_gen EXO_DEBUG_NO_GEN=1
run s0 EXO_STEP_STAGGER=0
