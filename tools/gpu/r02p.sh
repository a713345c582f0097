#!/bin/bash
# round 2, call p (8 GPUs): ceiling of the host-buffer path at N = 1/2/4/8, then the driver's multi-GPU bench command at N = 8 and 2
mkdir -p gpurun_out/r02p
nvidia-smi topo -m > gpurun_out/r02p/topo.txt 2>&1; nproc >> gpurun_out/r02p/topo.txt; numactl -H >> gpurun_out/r02p/topo.txt 2>&1
for N in 1 2 4 8; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/e2e_probe_multi.py --pin 1 2>/dev/null | grep "^N=" | tee -a gpurun_out/r02p/e2e_probe.txt
done
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tools/e2e_probe_multi.py --pin 1 --chunks 8 2>/dev/null | grep "^N=" | tee -a gpurun_out/r02p/e2e_probe.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 tools/e2e_probe_multi.py --pin 1 --chunks 2 2>/dev/null | grep "^N=" | tee -a gpurun_out/r02p/e2e_probe.txt
for N in 8 2; do
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02p/bench_gpus$N.json 2> gpurun_out/r02p/bench_gpus$N.err; echo "bench N=$N rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02p/bench_gpus$N.json').read().strip().splitlines()[-1])
    print('N=$N value %.3e ms_per_step %.4f frac %.3f e2e %.3e'%(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value']), d['e2e'].get('d2h_ceiling'))
    print(d.get('cfg4'))
except Exception as ex:
    print('parse failed', ex); print(open('gpurun_out/r02p/bench_gpus$N.err').read()[-1500:])
PY
done
