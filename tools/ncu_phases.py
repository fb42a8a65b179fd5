#!/usr/bin/env python
"""Split an .ncu-rep source page at BAR.SYNC instructions: per segment, warp-instructions executed and stall samples
by reason. usage: ncu_phases.py file.ncu-rep"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = None; segs = []; cur = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        if hdr is not None: break
        continue
    if hdr is None:
        if r and r[0] == 'Address':
            hdr = r; ia = hdr.index('Instructions Executed'); isrc = hdr.index('Source'); ismp = hdr.index('# Samples')
            stall_cols = [(i, c) for i, c in enumerate(hdr) if c.startswith('stall_') and '(Not Issued)' not in c]
            cur = {'n': 0, 'smp': 0, 'st': collections.Counter(), 'first': None, 'ops': collections.Counter()}
        continue
    ins = r[isrc].strip()
    if cur['first'] is None: cur['first'] = ins
    cur['n'] += int(r[ia]); cur['smp'] += int(r[ismp] or 0)
    for i, c in stall_cols: cur['st'][c] += int(r[i] or 0)
    op = ins.split(); o = op[0] if not op[0].startswith('@') else op[1]
    cur['ops'][o.split('.')[0]] += int(r[ia])
    if 'BAR.SYNC' in ins or ins.startswith('EXIT'):
        segs.append(cur); cur = {'n': 0, 'smp': 0, 'st': collections.Counter(), 'first': None, 'ops': collections.Counter()}
tot_s = sum(s['smp'] for s in segs) or 1; tot_n = sum(s['n'] for s in segs) or 1
for k, s in enumerate(segs):
    top = ', '.join(f"{c[6:]}={100*v/max(1,s['smp']):.0f}%" for c, v in s['st'].most_common(5))
    ops = ', '.join(f"{o}:{v*100//max(1,s['n'])}%" for o, v in s['ops'].most_common(6))
    print(f"seg {k}: instr {100*s['n']/tot_n:5.1f}%  samples {100*s['smp']/tot_s:5.1f}%  | {top} | {ops}")
