#!/bin/bash
# Experiment 1: MMA-issuer wait variants of the attention kernels (DDP_ATTN_DBG bits 2 / 4).
mkdir -p gpurun_out
L=gpurun_out/exp1.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $L 2>&1
for dbg in 0 2 4; do
  echo "=== DDP_ATTN_DBG=$dbg perfall" >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perfall >> $L 2>&1
  echo "exit $?" >> $L
done
for dbg in 0 2 4; do
  for mode in quick multi bwd; do
    echo "=== DDP_ATTN_DBG=$dbg $mode" >> $L
    DDP_ATTN_DBG=$dbg timeout 300 build/attn_harness $mode 2>&1 | grep -E "PASS|FAIL|SUMMARY|error|WATCHDOG" >> $L
    echo "exit ${PIPESTATUS[0]}" >> $L
  done
done
tail -5 $L
