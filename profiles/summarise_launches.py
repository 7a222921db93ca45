import csv,collections,sys
rows=list(csv.reader(open(sys.argv[1])))
for i,r in enumerate(rows):
    if 'Kernel Name' in r: h=i;break
hdr=rows[h]; ix={k:i for i,k in enumerate(hdr)}
d=collections.OrderedDict(); seq=[]
for r in rows[h+1:]:
    if len(r)<len(hdr): continue
    if r[ix['Metric Name']]!='gpu__time_duration.sum': continue
    name=r[ix['Kernel Name']].split('(')[0].split('::')[-1]
    v=float(r[ix['Metric Value']].replace(',','')); u=r[ix['Metric Unit']]
    if u in('ns','nsecond'): v/=1000
    elif u in('ms','msecond'): v*=1000
    elif u in('s','second'): v*=1e6
    seq.append((name,v)); d.setdefault(name,[]).append(v)
tot=sum(v for _,v in seq)
print('launches',len(seq),'total us',round(tot,1))
for k,v in sorted(d.items(), key=lambda kv:-sum(kv[1])):
    print(f'{k:40s} n={len(v):4d} avg={sum(v)/len(v):8.2f} us  share={sum(v)/tot*100:5.1f}%')
