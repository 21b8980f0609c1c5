import csv, sys
path = sys.argv[1]
rows=[r for r in csv.reader(open(path)) if r and r[0]!='kind']
n=len(rows)//3
rows=rows[:n]
tot={0:0.0,1:0.0}
for i,r in enumerate(rows):
    k,gf,mb,ms,tf=int(r[0]),float(r[1]),float(r[2]),float(r[3]),float(r[4])
    tot[k]+=ms
    print(f"{i:3d} {'igemm' if k==0 else 'wgrad'} gflop={gf:8.2f} minMB={mb:8.1f} ms={ms:7.4f} TF={tf:7.1f} GB/s(min)={mb/ms:8.1f}")
print(n,'launches/step', tot)
