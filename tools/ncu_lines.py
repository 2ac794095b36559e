"""Attribute an ncu SASS source page (ncu -i X.ncu-rep --page source --csv) to CUDA source lines using the line
table nvdisasm prints for the same cubin.

    python tools/ncu_lines.py sass.csv kernel.cubin [top_n]
"""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def main():
    sass_csv, cubin = sys.argv[1], sys.argv[2]
    top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout
    addr2line, cur, func = {}, None, None
    for ln in dis.splitlines():
        m = re.search(r"//## File \"([^\"]+)\", line (\d+)(.*)", ln)
        if m:
            inl = " (inlined)" if "inlined" in m.group(3) else ""
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            func = m.group(1)
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
        if m and cur:
            addr2line[(func, int(m.group(1), 16))] = cur
    rows = list(csv.reader(open(sass_csv)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hdr]
    i_s = h.index("# Samples")
    i_i = h.index("Instructions Executed")
    base = None
    agg = defaultdict(lambda: [0, 0])
    per_addr = []
    funcs = sorted({f for f, _ in addr2line})
    for r in rows[hdr + 1:]:
        if len(r) <= i_i:
            continue
        a = int(r[0], 16) if r[0].startswith("0x") else int(r[0])
        if base is None:
            base = a
        off = a - base
        s = int(float(r[i_s] or 0))
        n = int(float(r[i_i] or 0))
        key = None
        for f in funcs:
            if "latent" in f or len(funcs) == 1:
                key = addr2line.get((f, off))
                if key:
                    break
        agg[key][0] += s
        agg[key][1] += n
        per_addr.append((off, s, n, r[1]))
    tot_s = sum(v[0] for v in agg.values())
    tot_i = sum(v[1] for v in agg.values())
    print(f"total samples {tot_s}, warp instructions {tot_i}")
    for k, (s, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top_n]:
        print(f"{str(k):40s} samples {s:7d} ({100 * s / max(tot_s, 1):5.1f}%)  instr {n:9d} ({100 * n / max(tot_i, 1):5.1f}%)")


if __name__ == "__main__":
    main()
