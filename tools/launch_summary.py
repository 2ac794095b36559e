import csv, collections, re, sys
path=sys.argv[1]
lines=[l for l in open(path) if not l.startswith('==')]
rows=list(csv.DictReader(lines))
names=[r['Kernel Name'] for r in rows]
idx=[i for i,n in enumerate(names) if 'text_mask_select' in n]
s,e=(idx[-2],idx[-1]) if len(idx)>=2 else (0,len(rows))
agg=collections.defaultdict(lambda:[0,0.0]); tot=0
for r in rows[s:e]:
    n=re.sub(r'\(.*','',r['Kernel Name']); n=re.sub(r'^void ','',n)
    t=float(r['Metric Value'].replace(',',''))/1e3
    agg[n][0]+=1; agg[n][1]+=t; tot+=t
print(f'one step: {e-s} launches, {tot/1e3:.3f} ms of kernel time (ncu: cold caches, serialised)')
for n,(c,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:int(sys.argv[2]) if len(sys.argv)>2 else 22]:
    print(f'{t:10.1f} us {100*t/tot:5.1f}% x{c:4d} avg {t/c:7.1f} us  {n[:95]}')
