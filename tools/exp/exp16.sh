#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out/n${N}_final2
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29931 bench.py --gpus $N --steps 10 --warmup 3 2>&1 | grep -a '^{' > gpurun_out/n${N}_final2/bench_ours.json
python -c "
import json
d = json.load(open('gpurun_out/n${N}_final2/bench_ours.json')); print('device', d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks']); print(d['step_ms_rank0'])"
