#!/usr/bin/env python
"""Top warp-stall lines of one launch from `ncu -i X.ncu-rep --page source --csv` output.
usage: ncu_source_top.py file.csv [n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
print(rows[0][1][:110])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr) and r != hdr and (r[ix["# Samples"]] or "0").isdigit()]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
print("total samples", tot)
agg = {}
for s in stalls:
    agg[s] = sum(int(r[ix[s]] or 0) for r in body)
print("stall mix:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
body.sort(key=lambda r: -int(r[ix["# Samples"]] or 0))
for r in body[:n]:
    smp = int(r[ix["# Samples"]] or 0)
    top = sorted(((int(r[ix[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print("%5.1f%% %-8s %-70s %s" % (100.0 * smp / max(tot, 1), r[ix["Instructions Executed"]], r[ix["Source"]][:70],
                                    " ".join("%s:%d" % (b, a) for a, b in top if a)))
