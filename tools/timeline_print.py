import csv, re, sys
rows=list(csv.DictReader(open(sys.argv[1])))
for r in rows: r['start_us']=float(r['start_us']); r['dur_us']=float(r['dur_us'])
sel=[i for i,r in enumerate(rows) if 'adamw_kernel' in r['kernel']]
rows=rows[sel[0]+1:sel[1]+1]
t0=rows[0]['start_us']
lo=float(sys.argv[2]) if len(sys.argv)>2 else 0
for r in sorted(rows,key=lambda r:r['start_us']):
    k=re.sub(r'\(.*','',r['kernel'])[:70]
    if r['start_us']-t0>=lo: print(f"{r['start_us']-t0:8.1f} {r['dur_us']:7.1f} s{r['stream']:>4} {k}")
