import csv,sys,re,collections
rows=list(csv.reader(open(sys.argv[1])))
# find kernel sections: a row starting with "Kernel Name"
sections=[]; cur=None
for r in rows:
    if r and r[0]=="Kernel Name": cur={'name':r[1],'hdr':None,'rows':[]}; sections.append(cur); continue
    if cur is None: continue
    if cur['hdr'] is None: cur['hdr']=r; continue
    cur['rows'].append(r)
s=sections[0]; h=s['hdr']
ia=h.index("Instructions Executed"); isrc=h.index("Source"); iw=h.index("L1 Wavefronts Shared"); iwi=h.index("L1 Wavefronts Shared Ideal"); ismp=h.index("# Samples")
tot=0; hist=collections.Counter(); wv=collections.Counter(); wvi=collections.Counter()
for r in s['rows']:
    n=int(r[ia]); tot+=n
    op=r[isrc].split()
    o=op[0] if not op[0].startswith('@') else op[1]
    o=o.split('.')[0] if not o.startswith(('LDS','STS','LDG','STG')) else o
    hist[o]+=n; wv[o]+=int(r[iw] or 0); wvi[o]+=int(r[iwi] or 0)
print("total warp instr",tot, "per tile", tot/16384)
for o,n in hist.most_common(40): print(f"{o:22s} {n/16384:9.1f} /tile  {100*n/tot:5.1f}%   smem wf {wv[o]/16384:8.1f} ideal {wvi[o]/16384:8.1f}")
