"""Print the kernel sequence of one latent layer's backward from a bench.py --timeline CSV."""
import csv, re, sys
rows = list(csv.DictReader(open(sys.argv[1])))
emb = [i for i, r in enumerate(rows) if 'text_embed_fwd' in r['kernel']]
rows = rows[emb[0]:emb[1]]
def short(n):
    return re.sub(r'\(.*', '', n).replace('void ', '').replace('pio::', '')[:50]
dq = [i for i, r in enumerate(rows) if 'attn_bwd_dq' in r['kernel']]
a = dq[8]
for r in rows[a - 14:a + 14]:
    print(f"{r['stream']:>4} {float(r['start_us']):9.1f} {float(r['dur_us']):6.1f}  {short(r['kernel'])}")
