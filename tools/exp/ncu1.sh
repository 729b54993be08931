#!/bin/bash
# ncu --set full of the three second-generation attention kernels on the flagship per-GPU shape (harness prof mode)
set -x
build/attn_harness prof > gpurun_out/prof_plain.log 2>&1 || exit 1
for k in attn_fwd2_kernel attn_bwd2_rows_kernel attn_bwd2_cols_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/$k build/attn_harness prof > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
