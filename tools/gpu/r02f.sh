#!/bin/bash
# round 2, call f: generation passes on 1/2/4 side streams; block shapes again
mkdir -p gpurun_out/r02f
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02f/pytest.txt 2>&1; tail -2 gpurun_out/r02f/pytest.txt
B="python bench.py --steps 300 --warmup 20 --no-policy --no-lap --no-large --no-motor-full --cpu-steps 50 --e2e-steps 5"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r02f/$name.json 2> gpurun_out/r02f/$name.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r02f/$name.json').read().strip().splitlines()[-1])
print('$name', 'ms_per_step %.4f'%d['ms_per_step'], 'kernel %.4f'%d['roofline']['kernel_ms'], 'value %.3e'%d['value'])
PY
}
run streams4 A=1
run streams2 EXO_GEN_STREAMS=2
run streams1 EXO_GEN_STREAMS=1
run s4_warps4 EXO_GEN_WARPS=4
run s4_warps2 EXO_GEN_WARPS=2
run s4_grid148 EXO_GEN_GRID=148
run s4_grid296 EXO_GEN_GRID=296
run s4_grid592 EXO_GEN_GRID=592
