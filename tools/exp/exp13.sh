#!/bin/bash
# cols2 distributed epilogue (store warp) at N=2: correctness, then timing with ablations
build/attn_harness bwd 2>&1 | tail -1
build/attn_harness multi 2>&1 | tail -1
timeout 200 python -m pytest tests/test_gpu_multirank.py -x -q 2>&1 | tail -2
run() { DDP_ATTN_DBG=$1 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29901 tools/time_flash_ops.py 2>&1 | grep -a '^{' | sed "s/^/dbg=$1 /"; }
run 0
run 16384
run 4096
