#!/bin/bash
# BASELINE.json configs 2, 3 and 5 on N GPUs (default 1): this framework (benchmarks/benchmark.py) and the
# unmodified reference (baseline/bench_ops_reference.py).  JSON lines go to $OUT (default gpurun_out/configs.jsonl).
N=${1:-1}
OUT=${2:-gpurun_out/configs_n$N.jsonl}
MODE=${3:-full}   # "reduced" = the largest point of every sweep only (multi-GPU budget)
ARMS=${4:-both}   # "ours" = skip the reference arm
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p "$(dirname "$OUT")"
: > "$OUT"
port=29700
ours() { port=$((port+1)); timeout 300 $RUN --master-port $port benchmarks/benchmark.py --file /tmp/bench_$$.json --no-dense "$@" 2>/dev/null | grep '^{' >> "$OUT"; rm -f /tmp/bench_$$.json; }
ref()  { [ "$ARMS" = ours ] && return; port=$((port+1)); timeout 300 $RUN --master-port $port baseline/bench_ops_reference.py "$@" 2>/dev/null | grep '^{' >> "$OUT"; }
# config 2: nt sweep, d = 768, fp32 (tf32), fwd+bwd
TS="1024 4096 8192 16384 24576"; DS="2 16 64 256 768"
if [ "$MODE" = reduced ]; then TS="24576"; DS="768"; fi
for T in $TS; do
  ours --mode nt --length $T --dim 768 --dtype fp32 --backward --offset 1024
  ref  --mode nt --length $T --dim 768 --backward --offset 1024
done
# config 3: all / tn, T = 8192, d sweep, fwd+bwd
for D in $DS; do
  for M in all tn; do
    ours --mode $M --length 8192 --dim $D --dtype fp32 --backward --offset 1024
    ref  --mode $M --length 8192 --dim $D --backward --offset 1024
  done
done
# config 5: non-local block attention, T = 16*56*56 = 50176, d = 512, fwd+bwd
ours --mode attn --length 50176 --dim 512 --heads 1 --dtype bf16 --backward --offset 1024
ref  --mode attn --length 50176 --dim 512 --heads 1 --backward --offset 1024
cat "$OUT"
