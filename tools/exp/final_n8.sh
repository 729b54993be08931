#!/bin/bash
N=${1:-8}
OUT=gpurun_out/n${N}_final
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$RUN --master-port 29811 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_ours.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours.log | tail -1 > $OUT/bench_ours.json; cut -c1-200 $OUT/bench_ours.json
$RUN --master-port 29812 tools/time_flash_ops.py 2>&1 | grep -a "^{" > $OUT/time_flash_ops.json; cat $OUT/time_flash_ops.json
$RUN --master-port 29813 bench.py --gpus $N --steps 5 --warmup 3 --T 65536 > $OUT/bench_ours_T65536.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours_T65536.log | tail -1 > $OUT/bench_ours_T65536.json; cut -c1-200 $OUT/bench_ours_T65536.json
$RUN --master-port 29814 bench.py --gpus $N --steps 10 --warmup 3 --mask causal --mask-refresh 4 > $OUT/bench_ours_causal.log 2>&1; grep -a '^{"impl"' $OUT/bench_ours_causal.log | tail -1 > $OUT/bench_ours_causal.json; cut -c1-200 $OUT/bench_ours_causal.json
$RUN --master-port 29815 tools/heap_report.py > $OUT/heap_report.log 2>&1; grep -a "rank 0" $OUT/heap_report.log
