"""Summarise `ncu -i X.ncu-rep --page source --csv` output: stall totals and the top-N
instructions by sampled stalls.  usage: python profiles/ncu_src_top.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]
data = [r for r in rows[hi + 1:] if len(r) == len(h)]
si, src, ie = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
num = lambda s: int(s) if s.isdigit() else 0
print("total samples", sum(num(r[si]) for r in data), "instructions", sum(num(r[ie]) for r in data))
agg = {h[i]: sum(num(r[i]) for r in data) for i in cols}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0})
for r in sorted(data, key=lambda r: -num(r[si]))[:n]:
    st = {h[i][6:]: num(r[i]) for i in cols if num(r[i]) > 0}
    print(r[si].rjust(6), r[ie].rjust(8), r[src].strip()[:64].ljust(64), dict(sorted(st.items(), key=lambda kv: -kv[1])[:3]))
