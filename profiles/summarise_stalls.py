import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]
ix={h:i for i,h in enumerate(hdr)}
data=[]
for r in rows[2:]:
    if len(r)!=len(hdr): break
    if not r[ix['# Samples']].isdigit(): break
    data.append(r)
tot=sum(int(r[ix['# Samples']]) for r in data)
print('total samples',tot, 'insts', len(data))
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg={s:sum(int(r[ix[s]]) for r in data) for s in stalls}
print(sorted(agg.items(), key=lambda x:-x[1])[:10])
N=int(sys.argv[2]) if len(sys.argv)>2 else 45
top=sorted(range(len(data)),key=lambda i:-int(data[i][ix['# Samples']]))[:N]
for i in top:
    r=data[i]
    print(i, r[ix['# Samples']], r[1].strip()[:100], {s[6:]:r[ix[s]] for s in stalls if int(r[ix[s]])>0.2*int(r[ix['# Samples']])})
