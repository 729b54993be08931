#!/bin/bash
# Measurement set on N GPUs of one box.  usage: run_n.sh N [full|short]
N=${1:-1}; MODE=${2:-short}
OUT=gpurun_out/n$N
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29800
tr() { port=$((port+1)); timeout ${TMO:-400} $RUN --master-port $port "$@"; }
b() { name=$1; shift; tr bench.py --gpus $N "$@" > $OUT/$name.log 2>&1; grep -a '^{"impl"' $OUT/$name.log | tail -1 > $OUT/$name.json; cut -c1-220 $OUT/$name.json; }
rm -f $OUT/sweep_ours.jsonl $OUT/sweep_reference.jsonl
tr benchmarks/sweep.py --impl ours --out $OUT/sweep_ours.jsonl > $OUT/sweep_ours.log 2>&1; tail -2 $OUT/sweep_ours.log | cut -c1-200
tr benchmarks/sweep.py --impl reference --iters 5 --out $OUT/sweep_reference.jsonl > $OUT/sweep_reference.log 2>&1; tail -2 $OUT/sweep_reference.log | cut -c1-200
b bench_ours --steps 10 --warmup 3
b bench_reference --impl reference --steps 3 --warmup 3
b bench_ours_T65536 --steps 3 --warmup 3 --T 65536
if [ "$MODE" = full ]; then
  b bench_ours_causal --steps 10 --warmup 3 --mask causal --mask-refresh 4
  b bench_reference_causal --impl reference --steps 3 --warmup 3 --mask causal
  b bench_ours_fp32 --steps 5 --warmup 3 --dtype fp32
  b bench_reference_autocast --impl reference --steps 3 --warmup 3 --ref-autocast bf16
  if [ $N -gt 1 ]; then
    DDP_HEAP=ipc b bench_ours_heap_ipc --steps 10 --warmup 3
    DDP_MULTICAST=0 b bench_ours_heap_vmm_unicast --steps 10 --warmup 3
    DDP_HEAP=ipc tr tools/heap_report.py 2>&1 | grep -E "^rank 0" > $OUT/heap_report_ipc.log; cat $OUT/heap_report_ipc.log
    for m in 1 0; do DDP_MULTICAST=$m tr tools/heap_report.py 2>&1 | grep -E "^rank 0" ; done > $OUT/heap_report.log 2>&1; cat $OUT/heap_report.log
    python -m pytest tests/test_gpu_multirank.py -m gpu -q > $OUT/pytest_multirank.log 2>&1; tail -3 $OUT/pytest_multirank.log
  fi
fi
