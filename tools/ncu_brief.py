#!/usr/bin/env python
"""Brief of an .ncu-rep: key raw metrics + dynamic SASS mix. usage: ncu_brief.py file.ncu-rep [units_per_launch]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); h, u, r = rows[0], rows[1], rows[2]
d = dict(zip(h, r))
def g(k):
    try: return float(d[k].replace(',', ''))
    except Exception: return float('nan')
print('kernel', d.get('Kernel Name', '')[:110])
print('grid', d.get('launch__grid_size'), 'block', d.get('launch__block_size'), 'regs', d.get('launch__registers_per_thread'),
      'smem/blk', d.get('launch__shared_mem_per_block_dynamic'), 'occ limits regs/smem/warps', d.get('launch__occupancy_limit_registers'),
      d.get('launch__occupancy_limit_shared_mem'), d.get('launch__occupancy_limit_warps'))
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_lgds.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_active.avg',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum', 'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum']
for k in keys:
    for kk in h:
        if kk.endswith(k):
            v = d[kk]
            try: per = float(v.replace(',', '')) / units
            except Exception: per = float('nan')
            print(f'  {k:75s} {v:>18s} {dict(zip(h,u))[kk]:10s} per-unit {per:12.2f}')
            break
for k in h:
    if 'issue_stalled' in k and k.endswith('per_issue_active.ratio'):
        try:
            if float(d[k]) > 0.15: print(f'  stall {k.split("issue_stalled_")[1].split("_per_issue")[0]:28s} {float(d[k]):.3f}')
        except Exception: pass
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = None; hist = collections.Counter(); wv = collections.Counter(); wvi = collections.Counter(); tot = 0; done = False
for r in rows:
    if r and r[0] == 'Kernel Name':
        if hdr is not None: break
        continue
    if hdr is None:
        if r and r[0] == 'Address': hdr = r; ia = hdr.index('Instructions Executed'); isrc = hdr.index('Source'); iw = hdr.index('L1 Wavefronts Shared'); iwi = hdr.index('L1 Wavefronts Shared Ideal')
        continue
    n = int(r[ia]); tot += n
    op = r[isrc].split(); o = op[0] if not op[0].startswith('@') else op[1]
    o = o.split('.')[0] if not o.startswith(('LDS', 'STS', 'LDG', 'STG', 'LDL', 'STL', 'LDC')) else o
    hist[o] += n; wv[o] += int(r[iw] or 0); wvi[o] += int(r[iwi] or 0)
print(f'dynamic SASS: {tot} warp instr, per unit {tot/units:.1f}')
for o, n in hist.most_common(28):
    print(f'  {o:22s} {n/units:9.1f} /unit {100*n/tot:5.1f}%   smem wavefronts {wv[o]/units:8.1f} (ideal {wvi[o]/units:8.1f})')
