import re,csv,sys
from collections import defaultdict
sass, prof = sys.argv[1], sys.argv[2]
SRC = sys.argv[4] if len(sys.argv) > 4 else 'mpcb_step.cuh'
cur=None; inst_line=[]; ops=[]
for ln in open(sass):
    if '//## File' in ln:
        ms=re.findall(r'File "([^"]+)", line (\d+)',ln)
        pick=None
        for f,l in ms:
            if f.endswith(SRC): pick=int(l)
        cur=pick if pick else cur; continue
    m=re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)',ln)
    if m: inst_line.append(cur); ops.append(m.group(2).split('.')[0])
rows=list(csv.reader(open(prof)))
hdr=rows[1]; ix={h:i for i,h in enumerate(hdr)}
data=[r for r in rows[2:] if len(r)==len(hdr) and r[0]!='Address']
if len(data)!=len(inst_line) and len(data)%len(inst_line)==0: data=data[:len(inst_line)]   # several captures of the same kernel: first one
assert len(data)==len(inst_line),(len(data),len(inst_line))
src=open('/root/repo/mpc_for_av_at_intersection_b200/csrc/'+SRC).read().split('\n')
marks=[(1,'file head')]
for i,l in enumerate(src):
    m=re.search(r'// \[phase: (.+?)\]',l)
    if m: marks.append((i+1,m.group(1)))
if len(marks)<10: sys.exit("ncu_phase.py: the '// [phase: ...]' markers are missing from mpcb_step.cuh -- refusing to write a table")
marks.sort()
agg=defaultdict(lambda:[0,0,0,defaultdict(int)])
for l,op,r in zip(inst_line,ops,data):
    name='?'
    for start,n in marks:
        if l and l>=start: name=n
    g=agg[name]; ex=int(r[ix['Instructions Executed']]); g[0]+=int(r[ix['# Samples']]); g[1]+=ex; g[2]+=1; g[3][op]+=ex
ts=sum(a[0] for a in agg.values()); ti=sum(a[1] for a in agg.values())
NP=float(sys.argv[3]) if len(sys.argv)>3 else 4096
print('total exec/problem', ti/NP)
for start,n in marks:
    g=agg[n]
    top=sorted(g[3].items(), key=lambda kv:-kv[1])[:7]
    print(f"{start:4d} {n:24s} smp {g[0]/ts:.3f} exec {g[1]/NP:7.0f} ({g[1]/ti:.3f}) static {g[2]:5d}  "+' '.join(f"{k}:{v/NP:.0f}" for k,v in top))
