import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1])))
# file has sections per kernel: "Kernel Name" row, header row, then instructions
secs=[]; cur=None
for r in rows:
    if r and r[0]=='Kernel Name': cur={'name':r[1],'hdr':None,'rows':[]}; secs.append(cur); continue
    if cur is None: continue
    if cur['hdr'] is None: cur['hdr']=r; continue
    cur['rows'].append(r)
s=secs[int(sys.argv[2]) if len(sys.argv)>2 else 0]
hdr=s['hdr']; iS=hdr.index('Source'); iE=hdr.index('Instructions Executed')
inst=[r for r in s['rows'] if len(r)>iE and r[iE].strip().isdigit()]
tot=sum(int(r[iE]) for r in inst)
print(s['name'][:60],'static',len(inst),'dynamic warp instrs',tot)
c=collections.Counter()
for r in inst:
    t=r[iS].split()
    op=t[1] if t[0].startswith('@') else t[0]
    c[op.split('.')[0]]+=int(r[iE])
for op,n in c.most_common(28): print(f'{op:12s} {n:12d} {100*n/tot:5.1f}%')
