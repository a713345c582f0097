#!/bin/bash
# round 2, call S: the driver's 2-GPU launch of both arms on the final code
mkdir -p gpurun_out/r02S
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02S/bench_gpus2.json 2> gpurun_out/r02S/bench_gpus2.err; echo "rc $?"; tail -c 300 gpurun_out/r02S/bench_gpus2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02S/bench_ref_gpus2.json 2> gpurun_out/r02S/bench_ref_gpus2.err; echo "rc $?"; tail -c 200 gpurun_out/r02S/bench_ref_gpus2.json
