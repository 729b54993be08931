#!/usr/bin/env python
"""Hottest SASS instructions of an .ncu-rep (source page): samples, executed count, dominant stall reasons."""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
recs = []
tot = 0
for r in rows[2:]:
    if len(r) < len(hdr): continue
    try: n = int(r[ix['# Samples']])
    except ValueError: continue
    tot += n
    recs.append((n, r))
recs.sort(key=lambda x: -x[0])
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
print('total samples', tot)
for n, r in recs[:top]:
    st = sorted(((int(r[ix[c]] or 0), c[6:]) for c in stall_cols), reverse=True)[:3]
    print('%5.1f%% %-62s exec=%s  %s' % (100.0 * n / tot, r[ix['Source']][:62], r[ix['Instructions Executed']], ' '.join('%s:%d' % (c, v) for v, c in st if v)))
