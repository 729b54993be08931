#!/bin/bash
# round-2 closing run on N GPUs: dense-twin test, flagship bench (ours), per-kernel timings
N=${1:-8}
OUT=gpurun_out/n${N}_final2
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 python -m pytest tests/test_gpu_multirank.py -x -q -k "dense_twin and $N" > $OUT/pytest.log 2>&1; tail -2 $OUT/pytest.log
timeout 200 $RUN --master-port 29811 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_ours.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours.log | tail -1 > $OUT/bench_ours.json; cut -c1-260 $OUT/bench_ours.json
timeout 200 $RUN --master-port 29812 tools/time_flash_ops.py 2>&1 | grep -a "^{" > $OUT/time_flash_ops.json; cat $OUT/time_flash_ops.json
