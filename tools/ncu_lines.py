#!/usr/bin/env python
"""per-source-line view of an .ncu-rep (needs -lineinfo and --import-source on): samples, executed warp instructions and
the dominant stall reasons of the hottest lines of one kernel.   usage: ncu_lines.py rep kernel-regex [top]"""
import csv, subprocess, sys, collections

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + kern],
                     capture_output=True, text=True).stdout
agg = collections.OrderedDict()
fname, col, seen_fn = None, None, set()
first_fn = None
for r in csv.reader([l for l in out.splitlines() if l.startswith('"')]):
    if r[0] == 'File Path':
        fname = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        if first_fn is None: first_fn = r[1]
        cur_fn = r[1]; continue
    if r[0] == 'Line No':
        col = {}
        for i, n in enumerate(r):
            col.setdefault(n, i)
        src2 = [i for i, n in enumerate(r) if n == 'Source'][1]
        continue
    if col is None or len(r) <= col['# Samples'] or cur_fn != first_fn or not r[0].isdigit():
        continue
    cur = agg.setdefault((fname, int(r[0])), {'src': r[1].strip(), 'smp': 0.0, 'ins': 0.0, 'st': collections.Counter()})
    f = lambda x: float(x) if x.replace('.', '', 1).isdigit() else 0.0
    cur['smp'] += f(r[col['# Samples']]); cur['ins'] += f(r[col['Instructions Executed']])
    for n, i in col.items():
        if n.startswith('stall_') and 'Not Issued' not in n:
            cur['st'][n[6:]] += f(r[i])
tot = sum(v['smp'] for v in agg.values()); toti = sum(v['ins'] for v in agg.values())
print('%s: total samples %d, warp instructions %.0f' % (first_fn, tot, toti))
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1]['smp'])[:top]:
    st = ' '.join('%s:%d' % (n, c) for n, c in v['st'].most_common(3) if c)
    print('%5.1f%% smp %5.1f%% ins %s:%-4d %-60s %s' % (100 * v['smp'] / max(tot, 1), 100 * v['ins'] / max(toti, 1), f[:12], ln, v['src'][:60], st))
