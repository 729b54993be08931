#!/bin/bash
L=gpurun_out/exp10.log
: > $L
for dbg in 0 256 512 768; do
  echo "=== DDP_ATTN_DBG=$dbg" >> $L
  DDP_ATTN_DBG=$dbg timeout 120 build/attn_harness perfall 2>&1 | grep BWD >> $L
done
