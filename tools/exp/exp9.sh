#!/bin/bash
L=gpurun_out/exp9.log
: > $L
for dbg in 0 128; do
  echo "=== DDP_ATTN_DBG=$dbg" >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perfall >> $L 2>&1
done
DDP_ATTN_DBG=128 timeout 300 build/attn_harness quick 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
DDP_ATTN_DBG=128 timeout 300 build/attn_harness bwd 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
