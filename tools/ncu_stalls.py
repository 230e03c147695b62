#!/usr/bin/env python
"""Top stall locations from `ncu -i X.ncu-rep --page source --csv` output.  usage: ncu_stalls.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    try:
        n = int(r[ix["# Samples"]])
    except (ValueError, IndexError):
        continue
    data.append((n, r))
tot = sum(n for n, _ in data)
print("total samples", tot)
agg = {}
for n, r in data:
    for s in stalls:
        v = r[ix[s]]
        if v not in ("", "0"):
            agg[s] = agg.get(s, 0) + int(v)
print("by reason:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
data.sort(key=lambda x: -x[0])
for n, r in data[:top_n]:
    st = {s: int(r[ix[s]]) for s in stalls if r[ix[s]] not in ("", "0")}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print("%6d %5.1f%% %s  %-72s %s" % (n, 100.0 * n / tot, r[ix["Address"]][-5:], r[ix["Source"]][:72], top))
