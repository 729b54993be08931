#!/usr/bin/env python
"""Summarises .ncu-rep files (ncu -i ... --page raw --csv): pipe utilisation, issue slots, stall reasons."""
import csv, io, subprocess, sys
KEYS = [('gpu__time_duration.sum', 'duration'), ('sm__cycles_elapsed.max', 'SM elapsed cycles (max)'),
        ('smsp__inst_executed.sum', 'warp instructions executed'), ('sm__inst_issued.avg.pct_of_peak_sustained_active', 'issue slots busy %'),
        ('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'XU (MUFU) pipe %'),
        ('sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe (hmma) active %'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe active %'),
        ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'FMA pipe %'),
        ('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'ALU pipe %'),
        ('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'LSU pipe %'),
        ('sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active', 'uniform pipe %'),
        ('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'shared-memory wavefronts'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'shared-memory bank conflicts'),
        ('dram__bytes_read.sum', 'DRAM bytes read'), ('dram__bytes_write.sum', 'DRAM bytes written'),
        ('smsp__inst_executed_op_local_ld.sum', 'local loads (spills)'), ('smsp__inst_executed_op_local_st.sum', 'local stores (spills)'),
        ('launch__registers_per_thread', 'registers / thread (launch)'), ('launch__shared_mem_per_block_dynamic', 'dynamic smem / CTA')]
def load(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    return {h: (v, u) for h, u, v in zip(hdr, units, vals)}
reps = sys.argv[1:]
data = [load(r) for r in reps]
names = [d.get('Kernel Name', ('?', ''))[0].split('(')[0][-40:] for d in data]
print('| metric | ' + ' | '.join(names) + ' |')
print('|---|' + '---|' * len(names))
for k, label in KEYS:
    if any(k in d for d in data):
        print('| %s | ' % label + ' | '.join('%s %s' % d.get(k, ('-', '')) for d in data) + ' |')
print()
print('| stall (per issued instruction) | ' + ' | '.join(names) + ' |')
print('|---|' + '---|' * len(names))
stalls = sorted({k for d in data for k in d if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio') and 'not_issued' not in k})
for k in stalls:
    vals = [float(d.get(k, ('0', ''))[0] or 0) for d in data]
    if max(vals) >= 0.05:
        print('| %s | ' % k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')] + ' | '.join('%.2f' % v for v in vals) + ' |')
