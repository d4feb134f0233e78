import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1])))
cur=None; agg=collections.OrderedDict(); seen_kernel=0
for r in rows:
    if r and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if r and r[0]=='Function Name': continue
    if r and r[0]=='Line No': hdr=r; iE=hdr.index('Instructions Executed'); continue
    if cur and r and r[0].strip().isdigit() and len(r)>iE:
        key=(cur,int(r[0]),r[1].strip()[:90])
        if r[iE].strip().isdigit(): agg[key]=agg.get(key,0)+int(r[iE])
tot=sum(agg.values())
print('total',tot)
for (f,l,s),n in sorted(agg.items(), key=lambda kv:-kv[1])[:int(sys.argv[2]) if len(sys.argv)>2 else 45]:
    print(f'{100*n/tot:5.1f}% {n:11d} {f}:{l}: {s}')
