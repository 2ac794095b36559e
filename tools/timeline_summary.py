"""Summarise a bench.py --timeline CSV: per-stream busy time, idle gaps on the main stream, top kernels."""
import collections
import csv
import re
import sys

rows = list(csv.DictReader(open(sys.argv[1])))
for r in rows:
    r["start_us"] = float(r["start_us"]); r["dur_us"] = float(r["dur_us"])
sel = [i for i, r in enumerate(rows) if "text_mask_select" in r["kernel"]]
if len(sel) >= 2:
    rows = rows[sel[0]:sel[1]]
t0, t1 = rows[0]["start_us"], max(r["start_us"] + r["dur_us"] for r in rows)
print(f"one step: {len(rows)} device activities, wall {t1 - t0:.1f} us")
by = collections.defaultdict(list)
for r in rows:
    by[r["stream"]].append(r)
for s, rs in sorted(by.items(), key=lambda x: -len(x[1])):
    busy = sum(r["dur_us"] for r in rs)
    print(f"  stream {s}: {len(rs)} activities, busy {busy:.1f} us")
# union of busy intervals over all streams
iv = sorted((r["start_us"], r["start_us"] + r["dur_us"]) for r in rows)
cov, cur_s, cur_e = 0.0, *iv[0]
gaps = []
for a, b in iv[1:]:
    if a > cur_e:
        cov += cur_e - cur_s; gaps.append((a - cur_e, cur_e)); cur_s, cur_e = a, b
    else:
        cur_e = max(cur_e, b)
cov += cur_e - cur_s
print(f"  any-stream busy {cov:.1f} us, idle {t1 - t0 - cov:.1f} us in {len(gaps)} gaps (median {sorted(g for g, _ in gaps)[len(gaps)//2]:.2f} us)")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    n = re.sub(r"\(.*", "", r["kernel"]).replace("void ", "")
    agg[n][0] += 1; agg[n][1] += r["dur_us"]
tot = sum(v[1] for v in agg.values())
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print(f"{t:9.1f} us {100 * t / tot:5.1f}% x{c:4d} avg {t / c:6.1f} us  {n[:90]}")
