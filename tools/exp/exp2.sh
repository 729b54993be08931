#!/bin/bash
L=gpurun_out/exp2.log
: > $L
for a in "0 0" "8 0" "12 0"; do timeout 100 build/attn_harness trace $a 2>&1 | grep -E "PERF|TRACE-" >> $L; done
for dbg in 8 12; do
  echo "=== DDP_ATTN_DBG=$dbg perfall" >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perfall >> $L 2>&1
  for mode in quick multi bwd; do
    DDP_ATTN_DBG=$dbg timeout 300 build/attn_harness $mode 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" >> $L
  done
done
