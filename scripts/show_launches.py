import csv, collections, re, sys
rows=list(csv.reader(open(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/launches.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[hi]; data=rows[hi+1:]
ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
launches=[(r[ki], float(r[vi].replace(',',''))) for r in data if len(r)>vi]
names=[l[0] for l in launches]
adam=[i for i,n in enumerate(names) if 'adam_kernel' in n or 'sgd_kernel' in n]
step=launches[adam[-2]+1:adam[-1]+1]
agg=collections.OrderedDict()
for n,t in step:
    k=re.sub(r'\(.*','',n).split('::')[-1]
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=t
tot=sum(v[1] for v in agg.values())
print(f"{'kernel':40s} {'n':>4s} {'us':>9s} {'share':>6s}")
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print(f"{k:40s} {v[0]:4d} {v[1]/1e3:9.1f} {v[1]/tot*100:5.1f}%")
print('launches/step',len(step),'total us',round(tot/1e3,1))
