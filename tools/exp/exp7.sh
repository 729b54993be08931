#!/bin/bash
# v2 forward kernel (default) vs v1 (DDP_ATTN_DBG=64): correctness + timing
L=gpurun_out/exp7.log
: > $L
for dbg in 0 64; do
  echo "=== DDP_ATTN_DBG=$dbg" >> $L
  DDP_ATTN_DBG=$dbg timeout 300 build/attn_harness quick 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perf >> $L 2>&1
done
timeout 300 build/attn_harness multi 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
