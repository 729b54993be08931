#!/bin/bash
N=$1
OUT=gpurun_out/n${N}_final
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ $N -gt 1 ]; then
$RUN --master-port 29821 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_ours.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours.log | tail -1 > $OUT/bench_ours.json; cut -c1-200 $OUT/bench_ours.json
fi
$RUN --master-port 29823 bench.py --gpus $N --steps 3 --warmup 3 --T 65536 > $OUT/bench_ours_T65536.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours_T65536.log | tail -1 > $OUT/bench_ours_T65536.json; cut -c1-200 $OUT/bench_ours_T65536.json
