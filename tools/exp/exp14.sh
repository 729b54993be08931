#!/bin/bash
# gather reuse at N=2: correctness, then the step time with and without it
timeout 300 python -m pytest tests/test_gpu_multirank.py -x -q 2>&1 | tail -3
for r in 1 0; do
DDP_GATHER_REUSE=$r timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus 2 --steps 20 --warmup 5 2>&1 | grep -a '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('reuse=$r', d['ms_per_step'], d.get('e2e'), d.get('self_check'), d.get('gpu_launches'))"
done
