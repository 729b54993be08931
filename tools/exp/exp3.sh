#!/bin/bash
L=gpurun_out/exp3.log
: > $L
for dbg in 0 32; do
  echo "=== DDP_ATTN_DBG=$dbg perf" >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perf >> $L 2>&1
done
for mode in quick multi; do
  DDP_ATTN_DBG=32 timeout 300 build/attn_harness $mode 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" >> $L
done
for a in "0 0" "32 0"; do timeout 100 build/attn_harness_trace trace $a 2>&1 | grep -E "PERF|TRACE-" >> $L; done
