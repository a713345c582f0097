#!/bin/bash
# round-2 first GPU call: is a real pybullet importable on the GPU box? + baseline measurements of the round-1 code
mkdir -p gpurun_out/r02a
(python -c "import pybullet, pybullet_data, gym; print(pybullet.__version__)" ; echo "exit=$?"
 python -c "import pybullet" ; echo "exit=$?"
 python -c "import gym"; echo "exit=$?"
 python -c "import gymnasium"; echo "exit=$?"
 timeout 60 python -m pip download pybullet==3.2.5 -d /tmp/pb --no-deps; echo "pip exit=$?"
 python -m pip list 2>/dev/null | grep -i -E "bullet|gym|mujoco|dm_control|brax") > gpurun_out/r02a/pybullet_probe.txt 2>&1
nvidia-smi topo -m > gpurun_out/r02a/topo.txt 2>&1; nproc >> gpurun_out/r02a/topo.txt; lscpu | head -30 >> gpurun_out/r02a/topo.txt
timeout 300 python tools/reset_probe.py > gpurun_out/r02a/reset_probe.txt 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r02a/bench_20.json 2> gpurun_out/r02a/bench_20.err
tail -12 gpurun_out/r02a/reset_probe.txt
