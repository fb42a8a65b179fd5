#!/usr/bin/env python
"""Stall samples of an .ncu-rep's SASS page grouped into address regions separated by given SASS line numbers, or
the top-N instructions by samples.  usage: ncu_regions.py file.ncu-rep [b1,b2,...]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
bounds = [int(x) for x in sys.argv[2].split(',')] if len(sys.argv) > 2 else []
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; body = rows[2:]
col = {k: i for i, k in enumerate(hdr)}
stalls = [k for k in hdr if k.startswith('stall_') and 'Not Issued' not in k]
def region(i):
    r = 0
    for b in bounds:
        if i >= b: r += 1
    return r
agg = collections.defaultdict(collections.Counter)
for i, r in enumerate(body):
    g = region(i)
    agg[g]['samples'] += int(r[col['# Samples']] or 0)
    agg[g]['inst'] += int(r[col['Instructions Executed']] or 0)
    for s in stalls: agg[g][s] += int(r[col[s]] or 0)
for g in sorted(agg):
    a = agg[g]
    top = sorted(((a[s], s) for s in stalls), reverse=True)[:6]
    print(f'region {g}: inst {a["inst"]} samples {a["samples"]} samples/inst {a["samples"]/max(1,a["inst"])*1e3:.2f}e-3 ', ' '.join(f'{s[6:]}={n}' for n, s in top))
if not bounds:
    idx = sorted(range(len(body)), key=lambda i: -int(body[i][col['# Samples']] or 0))[:40]
    for i in sorted(idx):
        r = body[i]
        print(i, r[col['Source']].strip()[:70], r[col['# Samples']], r[col['Instructions Executed']])
