#!/bin/bash
# round 2, call Y: the driver's 8-GPU launch on the final code
mkdir -p gpurun_out/r02Y
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02Y/bench_gpus8.json 2> gpurun_out/r02Y/bench_gpus8.err; echo "rc $?"; tail -c 200 gpurun_out/r02Y/bench_gpus8.json
