#!/bin/bash
# round 2, call e: ncu --set full of the generation kernel (all 65 536 envs) and of the step kernel (final round-2 layout)
mkdir -p gpurun_out/r02e
python tools/ncu_target.py 65536 4 > gpurun_out/r02e/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:exo_gen_kernel -c 1 -o gpurun_out/r02e/gen python tools/ncu_target.py 65536 4 > gpurun_out/r02e/ncu_gen.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:exo_step_kernel -s 2 -c 1 -o gpurun_out/r02e/step python tools/ncu_target.py 65536 4 > gpurun_out/r02e/ncu_step.log 2>&1
tail -3 gpurun_out/r02e/ncu_gen.log gpurun_out/r02e/ncu_step.log
ls -la gpurun_out/r02e
