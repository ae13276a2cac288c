import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
tot={}
for r in rows[hdr+1:]:
    name=r[4].split('(')[0].replace('void ','').replace('kc::',''); t=float(r[-1])/1e3
    if len(sys.argv)>2: print(r[0], name, r[8], "%.1f us"%t)
    tot[name]=tot.get(name,0)+t
print({k:round(v,1) for k,v in tot.items()})
