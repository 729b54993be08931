#!/bin/bash
# device-timed vs end-to-end step at N=2, with and without gather reuse
for r in 1 0; do
DDP_GATHER_REUSE=$r timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2992$r bench.py --gpus 2 --steps 10 --warmup 3 2>&1 | grep -a '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('reuse=$r device', d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks'])"
done
