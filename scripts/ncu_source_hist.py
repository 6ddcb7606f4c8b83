#!/usr/bin/env python
"""opcode histogram + stall breakdown from an `ncu --page source --csv` export: scripts/ncu_source_hist.py file.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
def f(r, k):
    try: return float(r[idx[k]])
    except Exception: return 0.0
ti = sum(f(r, 'Instructions Executed') for r in data); ts = sum(f(r, '# Samples') for r in data)
print("warp instructions %.0f  samples %.0f  sass lines %d" % (ti, ts, len(data)))
op = collections.Counter(); ops = collections.Counter()
for r in data:
    s = r[idx['Source']].split()
    o = (s[1] if s and s[0].startswith('@') else (s[0] if s else '')).split('.')[0]
    op[o] += f(r, 'Instructions Executed'); ops[o] += f(r, '# Samples')
for o, c in op.most_common(22): print("%-10s inst %.3f  samples %.3f" % (o, c / ti, ops[o] / ts))
for k in ['stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_math', 'stall_wait', 'stall_not_selected', 'stall_selected',
          'stall_lg', 'stall_dispatch', 'stall_branch_resolving', 'stall_no_inst', 'stall_barrier']:
    print("%-24s %.3f" % (k, sum(f(r, k) for r in data) / ts))
print("shared wavefronts %.0f ideal %.0f" % (sum(f(r, 'L1 Wavefronts Shared') for r in data), sum(f(r, 'L1 Wavefronts Shared Ideal') for r in data)))
