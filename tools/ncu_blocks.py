#!/usr/bin/env python
"""Basic-block view of an ncu SASS-level source page:  ncu -i rep --page source --csv --print-source sass > x.csv ; python tools/ncu_blocks.py x.csv <section> [min share]
Groups consecutive SASS instructions with equal execution counts and prints each block's share of the executed warp instructions,
of the stall samples, the average active threads and its opcode mix; then the stall-reason totals and the opcode histogram."""
import csv,sys
from collections import Counter
rows=list(csv.reader(open(sys.argv[1])))
which=int(sys.argv[2]); thr=float(sys.argv[3]) if len(sys.argv)>3 else 0.006
hdr=[i for i,r in enumerate(rows) if r and r[0]=='Address']
H=rows[hdr[which]]; ci={n:i for i,n in enumerate(H)}   # (each section has its own header: a kernel with local-memory traffic has one more column)
sec=rows[hdr[which]+1:(hdr[which+1]-1 if which+1<len(hdr) else len(rows))]
ins=[r for r in sec if len(r)>6 and r[5].isdigit()]
tot=sum(int(r[5]) for r in ins); tsmp=sum(int(r[2]) for r in ins)
print("total inst %.3e samples %d"%(tot,tsmp))
st=Counter()
for r in ins:
    for k in H:
        if k.startswith('stall_') and '(Not' not in k and r[ci[k]].isdigit(): st[k]+=int(r[ci[k]])
print([(k,v) for k,v in st.most_common(8)])
blocks=[];cur=None
for r in ins:
    s=r[1]; op=(s.split()[0] if not s.startswith('@') else s.split()[1]).split('.')[0]; c=int(r[5])
    if cur and cur['c']==c: cur['n']+=1; cur['ops'].append(op); cur['smp']+=int(r[2]); cur['thr']+=int(r[6])
    else: cur={'a':r[0],'c':c,'n':1,'ops':[op],'smp':int(r[2]),'thr':int(r[6])}; blocks.append(cur)
for b in blocks:
    w=b['c']*b['n']
    if w/tot>thr:
        oc=Counter(b['ops']).most_common(7)
        print(f"{b['a'][-5:]} n={b['n']:4d} exec={b['c']:.2e} inst%={100*w/tot:5.1f} smp%={100*b['smp']/tsmp:5.1f} thr={b['thr']/max(w,1):4.1f} {oc}")
hist=Counter()
for r in ins:
    s=r[1]; op=(s.split()[0] if not s.startswith('@') else s.split()[1]).split('.')[0]; hist[op]+=int(r[5])
print([(k,f"{100*v/tot:.1f}") for k,v in hist.most_common(22)])
